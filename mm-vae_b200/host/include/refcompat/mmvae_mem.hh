// Reference header name -> this build (include/mmvae_mem.hh: host memory blocks for the dense minibatch).  No dense
// minibatch exists on this path (the data stay CSR, resident on the device).
#ifndef MMVAE_B200_REFCOMPAT_MMVAE_MEM_HH_
#define MMVAE_B200_REFCOMPAT_MMVAE_MEM_HH_
#endif
