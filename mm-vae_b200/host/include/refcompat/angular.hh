// Reference header name -> this build (include/modules/angular.hh: the Angular layer).  On this path the layer is device
// code (csrc/angular.cu, selected by the vMF model); there is no host-side class to forward to.
#ifndef MMVAE_B200_REFCOMPAT_ANGULAR_HH_
#define MMVAE_B200_REFCOMPAT_ANGULAR_HH_
#include "../mmvae_model.hh"
#endif
