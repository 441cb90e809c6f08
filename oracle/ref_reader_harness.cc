// TEST INFRASTRUCTURE ONLY (oracle/). Never linked into or called by the product.
//
// Drives the REFERENCE's own bgzf Matrix-Market reader, compiled where it lies
// under /root/reference: include/mmutil_index.hh, include/mmutil_bgzf_util.hh,
// include/utils/strbuf.hh, src/ext/tabix/{bgzf,kstring}.c,
// src/utils/{gzstream,bgzstream}.cc.   Sub-commands dump what the product's
// own reader must reproduce bit for bit:
//   index  <mtx> <idx>                 build_mmutil_index     (mmutil_index.hh:138-190)
//   table  <idx> <out.i64>             read_mmutil_index      (mmutil_index.hh:192-228)
//   header <mtx>                       peek_bgzf_header       (mmutil_bgzf_util.hh:155-251)
//   block  <mtx> <beg> <end> <out>     visit_bgzf_block       (mmutil_bgzf_util.hh:53-151)
//   batch  <mtx> <idx> <cols.i64> <B> <out.f32>   dense B x D minibatch; restates
//                                      mtx_data_block_t::read (mmvae_io.hh:150-245),
//                                      which itself cannot compile here (needs Eigen)
//   floats <in.txt> <out.f32>          strbuf_t::take_float   (strbuf.hh:72-131) per token
//   ones   <N> <out.mtx.gz>            text of create_ones_like (mmvae_io.hh:292-310,
//                                      io.hh:189-242) through the reference obgzf_stream
#include "util.hh"
#include "io.hh"
#include "mmutil_bgzf_util.hh"
#include "mmutil_index.hh"

#include <cstdio>
#include <cstring>

using Index = std::ptrdiff_t;

struct triplet_dump_t {
    std::vector<int64_t> rows, cols;
    std::vector<float> vals;
    void set_file(BGZF *) { }
    void eval_after_header(Index, Index, Index) { }
    void eval(Index r, Index c, float w)
    {
        rows.push_back(r);
        cols.push_back(c);
        vals.push_back(w);
    }
    void eval_end_of_file() { }
};

template <typename T>
static void
write_vec(const std::string &path, const std::vector<T> &v)
{
    std::ofstream f(path, std::ios::binary);
    f.write((const char *)v.data(), v.size() * sizeof(T));
}

template <typename T>
static std::vector<T>
read_vec(const std::string &path)
{
    std::ifstream f(path, std::ios::binary | std::ios::ate);
    size_t n = f.tellg();
    f.seekg(0);
    std::vector<T> v(n / sizeof(T));
    f.read((char *)v.data(), n);
    return v;
}

// ---- restatement of mmvae_io.hh:150-204 (find_consecutive_blocks) ----------
struct mem_block_t { Index lb, lb_mem, ub, ub_mem; };

static std::vector<mem_block_t>
consecutive_blocks(const std::vector<Index> &tab, const std::vector<Index> &subcol,
                   const Index gap = 10)
{
    const Index N = tab.size();
    std::vector<Index> s(subcol);
    std::sort(s.begin(), s.end());                         // :162-164
    std::vector<std::pair<Index, Index>> iv;
    Index beg = s[0], end = s[0];                          // :167-168
    for (size_t jj = 1; jj < s.size(); ++jj) {
        if (s[jj] >= end + gap) {                          // :172
            iv.emplace_back(beg, end + 1);
            beg = end = s[jj];
        } else {
            end = s[jj];                                   // :177
        }
    }
    if (beg <= s.back()) iv.emplace_back(beg, end + 1);    // :181-183
    std::vector<mem_block_t> ret;
    for (auto &p : iv) {
        Index lb = p.first, ub = p.second, ub_mem = 0;     // :190
        if (lb >= N) continue;                             // :193
        if (ub < N) ub_mem = tab[ub];                      // :198-200
        ret.push_back({ lb, tab[lb], ub, ub_mem });
    }
    return ret;
}

// ---- restatement of mmvae_io.hh:100-147,208-245 (_mem_reader_t, read) ------
struct mem_reader_t {
    Index D;
    const std::unordered_map<Index, Index> &remap;
    const std::vector<std::vector<Index>> &dup;
    std::vector<float> &mem;
    void set_file(BGZF *) { }
    void eval_after_header(Index, Index, Index) { }
    void eval(Index r, Index col, float w)
    {
        if (remap.count(col) == 0) return;                 // :124-125
        for (Index j : dup.at(remap.at(col))) mem[j * D + r] = w; // :127-130
    }
    void eval_end_of_file() { }
};

int
main(int argc, char **argv)
{
    using namespace mmutil::bgzf;
    using namespace mmutil::index;
    if (argc < 2) return 1;
    const std::string cmd = argv[1];

    if (cmd == "index" && argc == 4)
        return build_mmutil_index(argv[2], argv[3]);

    if (cmd == "table" && argc == 4) {
        std::vector<Index> tab;
        int rc = read_mmutil_index(argv[2], tab);
        write_vec(argv[3], std::vector<int64_t>(tab.begin(), tab.end()));
        return rc;
    }

    if (cmd == "header" && argc == 3) {
        mm_info_reader_t info;
        int rc = peek_bgzf_header(argv[2], info);
        std::printf("%ld %ld %ld\n", (long)info.rows(), (long)info.cols(), (long)info.nnz());
        return rc;
    }

    if (cmd == "block" && argc == 6) {
        triplet_dump_t d;
        int rc = visit_bgzf_block(argv[2], (Index)std::stoll(argv[3]),
                                  (Index)std::stoll(argv[4]), d);
        write_vec(std::string(argv[5]) + ".rows.i64", d.rows);
        write_vec(std::string(argv[5]) + ".cols.i64", d.cols);
        write_vec(std::string(argv[5]) + ".vals.f32", d.vals);
        return rc;
    }

    if (cmd == "batch" && argc == 7) {
        mm_info_reader_t info;
        CHK(peek_bgzf_header(argv[2], info));
        const Index D = info.rows(), B = std::stol(argv[5]);
        std::vector<Index> tab;
        CHK(read_mmutil_index(argv[3], tab));
        auto cols64 = read_vec<int64_t>(argv[4]);
        const size_t nb = cols64.size() / B;
        std::vector<float> all;
        for (size_t b = 0; b < nb; ++b) {
            std::vector<Index> subcol(cols64.begin() + b * B, cols64.begin() + (b + 1) * B);
            std::vector<float> mem(D * B, 0.f);
            std::vector<std::vector<Index>> dup(B);
            std::vector<Index> sorted(subcol);
            std::sort(sorted.begin(), sorted.end());
            std::unordered_map<Index, Index> col2idx;
            for (auto cc : sorted)
                if (col2idx.count(cc) < 1) { Index k = col2idx.size(); col2idx[cc] = k; } // :221-227
            for (Index j = 0; j < B; ++j) dup[col2idx[subcol[j]]].push_back(j);           // :229-233
            for (auto &blk : consecutive_blocks(tab, subcol)) {                            // :235-244
                mem_reader_t rd { D, col2idx, dup, mem };
                CHK(visit_bgzf_block(argv[2], blk.lb_mem, blk.ub_mem, rd));
            }
            all.insert(all.end(), mem.begin(), mem.end());
        }
        write_vec(argv[6], all);
        return 0;
    }

    if (cmd == "floats" && argc == 4) {
        std::ifstream f(argv[2]);
        std::string tok;
        std::vector<float> out;
        strbuf_t sb;
        while (f >> tok) {
            sb.clear();
            for (char ch : tok) sb.add(ch);
            out.push_back(sb.take_float());
        }
        write_vec(argv[3], out);
        return 0;
    }

    if (cmd == "ones" && argc == 4) {
        const long N = std::stol(argv[2]);
        obgzf_stream ofs(argv[3], std::ios::out);               // io.hh:233-236
        ofs << "%%MatrixMarket matrix coordinate integer general" << std::endl; // io.hh:197
        ofs << 1 << " " << N << " " << N << std::endl;          // io.hh:198
        const float one = 1.f;
        for (long j = 0; j < N; ++j)                            // io.hh:205-214 (row-major branch)
            ofs << 1 << " " << (j + 1) << " " << one << std::endl;
        ofs.close();
        return 0;
    }

    std::fprintf(stderr, "bad command\n");
    return 1;
}
