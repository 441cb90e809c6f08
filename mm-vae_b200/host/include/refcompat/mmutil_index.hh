// Reference header name -> this build (include/mmutil_index.hh: mmutil::index::build_mmutil_index, read_mmutil_index,
// the column index over bgzf virtual offsets): the code lives in ../mmvae_io.hh.
#ifndef MMVAE_B200_REFCOMPAT_MMUTIL_INDEX_HH_
#define MMVAE_B200_REFCOMPAT_MMUTIL_INDEX_HH_
#include "../mmvae_io.hh"
#endif
