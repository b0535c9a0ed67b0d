/* TEST INFRASTRUCTURE ONLY -- hand-written minimal stand-in for the header the
 * reference generates from src/config_schema.lua (process_schema.rg).  Only the
 * members the mixture constructors read are declared
 * (src/Mixtures/ConstPropMix.inl:31-52, AirMix.hpp:319-332,
 *  MultiComponent.inl:438-460).  Layout follows config_schema.lua:57-66,172-292. */
#ifndef HTR_ORACLE_STUB_CONFIG_SCHEMA_H
#define HTR_ORACLE_STUB_CONFIG_SCHEMA_H
#include <stdbool.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef enum { ViscosityModel_Constant, ViscosityModel_PowerLaw, ViscosityModel_Sutherland } ViscosityModel_enum;
typedef enum { MixtureModel_ConstPropMix, MixtureModel_IsentropicMix, MixtureModel_AirMix, MixtureModel_CH41StMix,
               MixtureModel_CH4_30SpMix, MixtureModel_CH4_43SpIonsMix, MixtureModel_CH4_26SpIonsMix,
               MixtureModel_FFCM1Mix, MixtureModel_BoivinMix, MixtureModel_H2_UCSDMix } MixtureModel_enum;
struct Species { char Name[10]; double MolarFrac; };
struct Mixture { struct { unsigned length; struct Species values[20]; } Species; };
struct ViscosityModel {
   int type;
   union {
      struct { double Visc; } Constant;
      struct { double ViscRef; double TempRef; } PowerLaw;
      struct { double ViscRef; double TempRef; double SRef; } Sutherland;
   } u;
};
struct MultiCompRef { double LRef; double PRef; double TRef; struct Mixture XiRef; };
struct MixtureModel {
   int type;
   union {
      struct { double gasConstant; double gamma; double prandtl; struct ViscosityModel viscosityModel; } ConstPropMix;
      struct { double gasConstant; double gamma; } IsentropicMix;
      struct MultiCompRef AirMix, CH41StMix, CH4_30SpMix, CH4_43SpIonsMix, CH4_26SpIonsMix, FFCM1Mix, BoivinMix, H2_UCSDMix;
   } u;
};
struct Config { struct { struct MixtureModel mixture; } Flow; };
#ifdef __cplusplus
}
#endif
#endif
