// Reference header name -> this build (include/utils/std_util.hh): nothing of it is needed on this path beyond <algorithm>.
#ifndef MMVAE_B200_REFCOMPAT_STD_UTIL_HH_
#define MMVAE_B200_REFCOMPAT_STD_UTIL_HH_
#include <algorithm>
#include <iterator>
#include <numeric>
#include <vector>
#endif
