// Reference header name -> this build (include/mmutil_bgzf_util.hh: visit_bgzf_block, peek_bgzf_header): ../mmvae_io.hh
// over the own BGZF reader ../mmvae_bgzf.hh.
#ifndef MMVAE_B200_REFCOMPAT_MMUTIL_BGZF_UTIL_HH_
#define MMVAE_B200_REFCOMPAT_MMUTIL_BGZF_UTIL_HH_
#include "../mmvae_bgzf.hh"
#include "../mmvae_io.hh"
#endif
