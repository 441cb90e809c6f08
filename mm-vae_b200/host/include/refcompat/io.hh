// Reference header name -> this build (include/io.hh: write_vector_file at global scope, io.hh:318-331).
#ifndef MMVAE_B200_REFCOMPAT_IO_HH_
#define MMVAE_B200_REFCOMPAT_IO_HH_
#include "../mmvae_model.hh"
template <typename Vec>
inline void
write_vector_file(const std::string filename, const Vec &out)
{
    const std::vector<float> v(out.begin(), out.end());
    ASSERT(mmvae::write_vector_file(filename, v) == EXIT_SUCCESS, "unable to write " << filename);
}
#endif
