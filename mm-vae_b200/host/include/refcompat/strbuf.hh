// Reference header name -> this build (include/utils/strbuf.hh: take_float / take_uint64 token parsers): ../mmvae_io.hh.
#ifndef MMVAE_B200_REFCOMPAT_STRBUF_HH_
#define MMVAE_B200_REFCOMPAT_STRBUF_HH_
#include "../mmvae_io.hh"
#endif
