// Reference header name -> this build (include/mmvae_io.hh: mtx_data_block_t, create_ones_like, write_tensor).
#ifndef MMVAE_B200_REFCOMPAT_MMVAE_IO_HH_
#define MMVAE_B200_REFCOMPAT_MMVAE_IO_HH_
#include "../mmvae_io.hh"
// mmvae_io.hh:292-310 of the reference: create_ones_like(data_block, out_file), global scope, no return value
inline void
create_ones_like(const mmvae::mtx_data_block_t &data_block, const std::string out_file)
{
    ASSERT(mmvae::create_ones_like((int64_t)data_block.ntot(), out_file) == EXIT_SUCCESS, "unable to write " << out_file);
}
#endif
