// Compile-time mixture selection (one library per mixture, like the reference's -DEOS=<Mix>:
// src/prometeo_const.h:37-41,90-92).  Shared by host (.cpp) and device (.cu) code.
#pragma once
#include <stdint.h>

#if defined(HTR_MIX_ConstPropMix)
#  define HTR_MIX_NAME "ConstPropMix"
#  define HTR_NSPEC 1
#  define HTR_MULTICOMP 0
#elif defined(HTR_MIX_AirMix)
   // src/Mixtures/AirMix.hpp:35-52
#  define HTR_MIX_NAME "AirMix"
#  define HTR_NSPEC 5
#  define HTR_MULTICOMP 1
#  define HTR_NREAC 2
#  define HTR_NTBREAC 3
#  define HTR_NFOREAC 0
#  define HTR_MAX_REACTANTS 2
#  define HTR_MAX_PRODUCTS 2
#  define HTR_MAX_TB 5
#elif defined(HTR_MIX_CH41StMix)
   // src/Mixtures/CH41StMix.hpp:34-53: one-step methane oxidation with specified forward orders (FWD_ORDERS)
#  define HTR_MIX_NAME "CH41StMix"
#  define HTR_NSPEC 4
#  define HTR_MULTICOMP 1
#  define HTR_NREAC 1
#  define HTR_NTBREAC 0
#  define HTR_NFOREAC 0
#  define HTR_MAX_REACTANTS 2
#  define HTR_MAX_PRODUCTS 2
#  define HTR_MAX_TB 0
#  define HTR_FWD_ORDERS 1
#else
#  error "define HTR_MIX_<Mixture> (ConstPropMix | AirMix | CH41StMix)"
#endif

#define HTR_NEQ (HTR_NSPEC + 4)

namespace htr {
constexpr int nSpec = HTR_NSPEC;
constexpr int nEq = HTR_NEQ;
constexpr int irU = nSpec;       // src/prometeo_const.h:90-92
constexpr int irE = nSpec + 3;

// src/Utils/constants.h:36-50
constexpr double PI = 3.1415926535898;
constexpr double RGAS = 8.3144598;
constexpr double Na = 6.02214086e23;
constexpr double kb = 1.38064852e-23;
constexpr double eps_0 = 8.8541878128e-12;
constexpr double eCrg = 1.60217662e-19;
}
