// Reference header name -> this build (include/utils/util.hh: TLOG / ELOG / WLOG / CHK / ASSERT, file_exists, zeropad).
#ifndef MMVAE_B200_REFCOMPAT_UTIL_HH_
#define MMVAE_B200_REFCOMPAT_UTIL_HH_
#include "../mmvae_util.hh"
#endif
