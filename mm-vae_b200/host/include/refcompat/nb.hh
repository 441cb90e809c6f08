// Reference header name -> this build (include/models/nb.hh: nbvae_options_t + parser, the strong dimension types, nbvae_t,
// nbvae_recorder_t, mmvae::nb::loss).  The model itself is csrc/ behind the C ABI.
#ifndef MMVAE_B200_REFCOMPAT_NB_HH_
#define MMVAE_B200_REFCOMPAT_NB_HH_
#include "../mmvae_options.hh"
#include "../mmvae_model.hh"
#endif
