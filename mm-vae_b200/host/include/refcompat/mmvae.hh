// Reference header name -> this build (include/mmvae.hh: mmvae_options_t, parse_mmvae_options).  Part of
// host/include/refcompat: put this directory BEFORE host/include on the include path and a translation unit written
// against the reference (its two mains, src/nb_vae_main.cc / src/vmf_vae_main.cc) compiles unchanged.
#ifndef MMVAE_B200_REFCOMPAT_MMVAE_HH_
#define MMVAE_B200_REFCOMPAT_MMVAE_HH_
#include "../mmvae_options.hh"
#include "../mmvae_model.hh"
// The reference's LOSS functors are declared to return torch::Tensor.  No libtorch on this path: the name resolves to
// what mmvae::nb::loss / mmvae::vmf::vmf_vae_loss hand back here (the KL weight; see mmvae_model.hh loss_request_t).
#ifndef MMVAE_B200_NO_TORCH_SHIM
namespace torch {
using Tensor = ::mmvae::loss_request_t;
}
#endif
#endif
