// TEST INFRASTRUCTURE ONLY (oracle/). Not part of the product path.
//
// Minimal stand-ins that let the reference's model headers
// (/root/reference/include/models/{nb,vmf}.hh, operators.hh, modules/angular.hh)
// compile verbatim against libtorch 2.11 without Eigen.  Only the *recorder*
// templates in nb.hh:569-662 / vmf.hh:457-551 touch these names; the harness
// never instantiates the recorders, the declarations merely have to parse.
#ifndef ORACLE_REF_SHIM_MODEL_SHIM_HH_
#define ORACLE_REF_SHIM_MODEL_SHIM_HH_

#include <string>
#include <cstdint>

namespace Eigen {
constexpr int Dynamic = -1;
constexpr int RowMajor = 1;
constexpr int ColMajor = 0;

template <typename S, int R, int C, int O = ColMajor>
struct Matrix {
    struct row_proxy {
        template <typename T>
        row_proxy &operator=(const T &) { return *this; }
    };
    int64_t rows() const { return 0; }
    int64_t cols() const { return 0; }
    void resize(int64_t, int64_t) { }
    void setZero() { }
    row_proxy row(int64_t) const { return row_proxy {}; }
};

template <typename M>
struct Map : M {
    template <typename... A>
    Map(A...) { }
};
} // namespace Eigen

// io.hh:561 and mmvae_io.hh:11 — declared only, never called by the harness.
template <typename T>
inline void
write_data_file(const std::string, const T &)
{
}

#endif
