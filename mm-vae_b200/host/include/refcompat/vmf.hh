// Reference header name -> this build (include/models/vmf.hh: vmf_options_t + parser, the strong dimension / scalar types,
// vmf_vae_t, vmf_vae_recorder_t, mmvae::vmf::vmf_vae_loss).  The model itself is csrc/ behind the C ABI.
#ifndef MMVAE_B200_REFCOMPAT_VMF_HH_
#define MMVAE_B200_REFCOMPAT_VMF_HH_
#include "../mmvae_options.hh"
#include "../mmvae_model.hh"
#endif
