// POD layouts of the mixture parameter block.  They are BYTE-IDENTICAL to the reference's
// `struct Mix` for the same mixture (ConstPropMix: src/Mixtures/ConstPropMix.hpp:80-93;
// multi-component: src/Mixtures/MultiComponent.hpp:109-143, Species.hpp:86-118,
// Reaction.hpp:100-236), because the Legion adapter receives that struct by value in every
// task's Args (src/prometeo_rhs.hpp:183) and forwards it to htr_set_mixture unchanged.
// tests/test_mixture.py compares these bytes with the reference's.
#pragma once
#include "htr_config.h"

namespace htr {

#if !HTR_MULTICOMP
struct MixData {
   double R;
   double gamma;
   int    viscosityModel;          // 0 Constant, 1 PowerLaw, 2 Sutherland
   double constantVisc;
   double powerlawTempRef;
   double powerlawViscRef;
   double sutherlandSRef;
   double sutherlandTempRef;
   double sutherlandViscRef;
   double Prandtl;
};
static_assert(sizeof(MixData) == 80, "ConstPropMix layout");
#else
struct NasaPoly {                  // three-range NASA-9 fit; switch temperatures are single precision
   float  TSwitch1, TSwitch2, TMin, TMax;
   double cpH[9], cpM[9], cpL[9];
};
struct Collision {                 // Lennard-Jones / Stockmayer parameters
   uint8_t Geom;                   // 0 atom, 1 linear, 3 non-linear
   double sigma, kbOveps, mu, alpha, Z298;
};
struct SpeciesData {
   char    Name[10];
   double  W;
   uint8_t inx;
   NasaPoly  cp;
   Collision diff;
};
static_assert(sizeof(SpeciesData) == 312, "species layout");

struct StoichTerm { uint8_t ind; float nu; };     // product (nu) or collider (efficiency)
#ifdef HTR_FWD_ORDERS
struct ReactTerm { uint8_t ind; float nu; float ord; };    // reactant with its specified order (src/Mixtures/Reaction.hpp:100-107)
#else
typedef StoichTerm ReactTerm;
#endif
struct Arrhenius  { double A; float n; float EovR; };
struct PlainReaction {
   Arrhenius k;
   bool    has_backward;
   uint8_t Neducts, Npducts;
   ReactTerm  educts[HTR_MAX_REACTANTS];
   StoichTerm pducts[HTR_MAX_PRODUCTS];
};
struct CollisionReaction {
   Arrhenius k;
   bool    has_backward;
   uint8_t Neducts, Npducts, Nthirdb;
   ReactTerm  educts[HTR_MAX_REACTANTS];
   StoichTerm pducts[HTR_MAX_PRODUCTS];
   StoichTerm thirdb[HTR_MAX_TB > 0 ? HTR_MAX_TB : 1];
};

struct MixData {
   SpeciesData species[HTR_NSPEC];
#if HTR_NREAC > 0
   PlainReaction reactions[HTR_NREAC];
#endif
#if HTR_NTBREAC > 0
   CollisionReaction tbReactions[HTR_NTBREAC];
#endif
   double PRef, TRef;
   double XiRef[HTR_NSPEC];
   double MixWRef, iMixWRef, rhoRef, ieRef, iuRef, iCpRef, imuRef, ilamRef, iDiRef, iwiRef, iKiRef, Eps0;
   double TMax, TMin;
};
#if defined(HTR_MIX_AirMix)
static_assert(sizeof(PlainReaction) == 56 && sizeof(CollisionReaction) == 96, "reaction layout");
static_assert(sizeof(MixData) == 2128, "AirMix layout (reference sizeof(Mix) = 2128)");
#endif
#endif

} // namespace htr
