// bgzf-indexed Matrix Market reader and the minibatch data block, B200 edition.
//
// Same surface and file formats as the reference:
//   mmutil::index::build_mmutil_index / read_mmutil_index   include/mmutil_index.hh:38-107,138-228
//   visit_bgzf_block, peek_bgzf_header                      include/mmutil_bgzf_util.hh:53-251
//   mmvae::mtx_data_block_t (read / clear / ntot / size / nfeature)   include/mmvae_io.hh:30-256
// What changes is what a batch IS.  The reference re-inflates and re-parses text for every batch
// and densifies it into a B x D float buffer; here the file is parsed ONCE into a cells x genes CSR
// (rows sorted by gene, last duplicate wins like mmvae_io.hh:127-130) that lives in HBM for the whole
// run, and read(batch) only records the cell ids handed to mmv_step.  The index-driven block path
// (find_consecutive_blocks + visit_bgzf_block) is kept verbatim in behaviour — dense_batch() uses it —
// so the bit-exact parity tests compare it with the reference reader's golden output.
#ifndef MMVAE_B200_IO_HH_
#define MMVAE_B200_IO_HH_

#include "mmvae_bgzf.hh"
#include "mmvae_util.hh"

#include <algorithm>
#include <chrono>
#include <tuple>
#include <unordered_map>
#include <cstdlib>
#include <functional>
#include <sys/stat.h>
#include <thread>

namespace mmutil {
namespace index {

using Index = std::ptrdiff_t;

struct mtx_header_t {
    Index max_row = 0, max_col = 0, max_elem = 0;
};

// first non-comment line with three fields: rows cols nnz  (mmutil_bgzf_util.hh:155-251)
inline int
peek_bgzf_header(const std::string &mtx_file, mtx_header_t &h)
{
    mmvae::bgzf_reader_t fp;
    ERR_RET(!fp.open(mtx_file), "unable to open " << mtx_file);
    std::string line, tok[3];
    while (true) {
        ERR_RET(fp.getline(line) < 0, "Failed to read the header");
        if (line.empty() || line[0] == '%') continue;
        if (mmvae::split_triplet_line(line, tok) == 3) {
            h.max_row = mmvae::take_uint64(tok[0].data(), tok[0].size());
            h.max_col = mmvae::take_uint64(tok[1].data(), tok[1].size());
            h.max_elem = mmvae::take_uint64(tok[2].data(), tok[2].size());
            return EXIT_SUCCESS;
        }
    }
}

// FUN: void(Index row0, Index col0, float weight).  Lines with < 3 fields and '%' comments are skipped
// (mmutil_bgzf_util.hh:104-109,129-136); indices are 1-based in the file (:139).  end_pos = 0 reads to EOF.
template <typename FUN>
int
visit_bgzf_block(const std::string &mtx_file, const int64_t beg_pos, const int64_t end_pos, FUN &&fun)
{
    mmvae::bgzf_reader_t fp;
    ERR_RET(!fp.open(mtx_file), "unable to open " << mtx_file);
    fp.seek(beg_pos);
    std::string line, tok[3];
    auto is_space = [](char c) { return c == ' ' || (c >= 9 && c <= 13); };
    auto is_digit = [](char c) { return c >= '0' && c <= '9'; };
    while (fp.getline(line) >= 0) {
        if (line.empty() || line[0] == '%') continue;
        // the shape every writer produces — "digits SP digits SP token", nothing else on the line — is split in
        // place (same three tokens the state machine below would cut, no copies); any other line takes the
        // reference's whitespace state machine
        const char *p = line.data(), *const e = p + line.size();
        const char *t0 = p;
        while (p < e && is_digit(*p)) ++p;
        const char *t0e = p, *t1 = nullptr, *t1e = nullptr, *t2 = nullptr;
        if (p > t0 && p < e && *p == ' ') {
            t1 = ++p;
            while (p < e && is_digit(*p)) ++p;
            t1e = p;
            if (p > t1 && p < e && *p == ' ') {
                t2 = ++p;
                while (p < e && !is_space(*p)) ++p;
                if (p != e || p == t2) t2 = nullptr;
            }
        }
        if (t2) {
            fun((Index)mmvae::take_uint64(t0, (size_t)(t0e - t0)) - 1, (Index)mmvae::take_uint64(t1, (size_t)(t1e - t1)) - 1,
                mmvae::take_float(t2, (size_t)(e - t2)));
        } else {
            if (mmvae::split_triplet_line(line, tok) < 3) continue;
            fun((Index)mmvae::take_uint64(tok[0].data(), tok[0].size()) - 1,
                (Index)mmvae::take_uint64(tok[1].data(), tok[1].size()) - 1, mmvae::take_float(tok[2].data(), tok[2].size()));
        }
        if (end_pos != 0 && fp.tell() >= end_pos) break;
    }
    return EXIT_SUCCESS;
}

// One scan: (column, virtual offset of the first line of that column), in file order
// (mmutil_index.hh:38-107); fails on unsorted columns (:76) and unindexed trailing columns (:171-179).
inline int
scan_mmutil_index(const std::string &mtx_file, std::vector<std::pair<Index, int64_t>> &col2file)
{
    ERR_RET(!mmvae::bgzf_reader_t::is_bgzf(mtx_file), "This file is not bgzipped: " << mtx_file);
    mmvae::bgzf_reader_t fp;
    ERR_RET(!fp.open(mtx_file), "unable to open " << mtx_file);
    std::string line, tok[3];
    Index max_col = -1;
    while (max_col < 0) {
        ERR_RET(fp.getline(line) < 0, "Failed to read the header");
        if (line.empty() || line[0] == '%') continue;
        if (mmvae::split_triplet_line(line, tok) == 3) max_col = mmvae::take_uint64(tok[1].data(), tok[1].size());
    }
    int64_t first_off = fp.tell(), last_off = first_off;
    col2file.clear();
    Index lineno = 0, last_col = 0;
    while (fp.getline(line) >= 0) {
        if (line.empty() || line[0] == '%') continue;
        if (mmvae::split_triplet_line(line, tok) < 3) continue;
        const Index col = (Index)mmvae::take_uint64(tok[1].data(), tok[1].size()) - 1;
        if (lineno == 0) {
            last_col = col;
            col2file.emplace_back(col, first_off);
        }
        if (col != last_col) {
            ERR_RET(!(col > last_col), "MTX must be sorted by columns");
            col2file.emplace_back(col, last_off);
            last_col = col;
        }
        last_off = fp.tell();
        ++lineno;
    }
    const Index last = col2file.empty() ? 0 : col2file.back().first;
    ERR_RET(last != max_col - 1, "Failed to index all the columns: " << last << " < " << max_col);
    return EXIT_SUCCESS;
}

// ${mtx}.index: plain gzip text, "col voffset\n" per distinct column (mmutil_index.hh:183-185, io.hh:248-266)
inline int
build_mmutil_index(const std::string &mtx_file, const std::string &index_file)
{
    std::vector<std::pair<Index, int64_t>> col2file;
    CHK(scan_mmutil_index(mtx_file, col2file));
    gzFile g = gzopen(index_file.c_str(), "wb");
    ERR_RET(!g, "unable to write " << index_file);
    for (auto &p : col2file) gzprintf(g, "%ld %ld\n", (long)p.first, (long)p.second);
    gzclose(g);
    TLOG("Created an index file: " << index_file);
    return EXIT_SUCCESS;
}

// int64[MaxIdx + 1] with the reference's gap fill (mmutil_index.hh:192-228): empty interior columns
// inherit the next column's offset; the loop stops at MaxIdx - 1.
inline int
read_mmutil_index(const std::string &index_file, std::vector<Index> &tab)
{
    gzFile g = gzopen(index_file.c_str(), "rb");
    ERR_RET(!g, "unable to read " << index_file);
    std::string text;
    char buf[1 << 16];
    int n;
    while ((n = gzread(g, buf, sizeof(buf))) > 0) text.append(buf, (size_t)n);
    gzclose(g);
    std::vector<std::pair<Index, Index>> pairs;
    std::stringstream ss(text);
    long a, b;
    while (ss >> a >> b) pairs.emplace_back((Index)a, (Index)b);
    ERR_RET(pairs.empty(), "empty file: " << index_file);
    Index max_idx = 0;
    for (auto &p : pairs) max_idx = std::max(max_idx, p.first);
    tab.assign((size_t)max_idx + 1, (Index)mmvae::MISSING_POS);
    for (auto &p : pairs) tab[(size_t)p.first] = p.second;
    for (Index j = 0; j < max_idx - 1; ++j)
        if (tab[(size_t)j] == mmvae::MISSING_POS) tab[(size_t)j] = tab[(size_t)j + 1];
    return EXIT_SUCCESS;
}

} // namespace index
} // namespace mmutil

namespace mmvae {

using Index = std::ptrdiff_t;

struct memory_block_t {   // mmvae_io.hh:30-35
    Index lb, lb_mem, ub, ub_mem;
};

// sorted batch columns -> merged [lb, ub) runs (gap 10) -> bgzf virtual-offset ranges (mmvae_io.hh:150-204)
inline std::vector<memory_block_t>
find_consecutive_blocks(const std::vector<Index> &index_tab, const std::vector<Index> &subcol, const Index gap = 10)
{
    const Index N = (Index)index_tab.size();
    std::vector<Index> s(subcol);
    std::sort(s.begin(), s.end());
    std::vector<std::pair<Index, Index>> iv;
    Index beg = s[0], end = s[0];
    for (size_t i = 1; i < s.size(); ++i) {
        if (s[i] >= end + gap) {
            iv.emplace_back(beg, end + 1);
            beg = end = s[i];
        } else {
            end = s[i];
        }
    }
    if (beg <= s.back()) iv.emplace_back(beg, end + 1);
    std::vector<memory_block_t> out;
    for (auto &b : iv) {
        if (b.first >= N) continue;
        const Index ub_mem = b.second < N ? index_tab[(size_t)b.second] : 0;
        out.push_back(memory_block_t{ b.first, index_tab[(size_t)b.first], b.second, ub_mem });
    }
    return out;
}

class mtx_data_block_t {
public:
    using Index = std::ptrdiff_t;

    // Reads the header and the index, then parses the whole file once into CSR.
    // csr_cache (or MMV_CSR_CACHE=1): keep the parsed CSR as ${idx_file}.csr beside the index, so that a later run
    // on the same, unchanged mtx file maps three arrays instead of inflating and parsing the text again.
    //
    // [col_lo, col_hi) (default: everything) makes this block ONE SHARD of the data points: only that range of the
    // text is inflated and parsed (the index table gives its bgzf virtual offsets) and the CSR holds hi - lo cells
    // with shard-local ids 0 .. hi - lo - 1.  The multi-GPU mains build one such block per GPU, so neither the
    // host nor any single device ever holds the whole matrix.  ntot() stays the GLOBAL count.
    mtx_data_block_t(const std::string &mtx_file, const std::string &idx_file, const Index batch_size, bool csr_cache = false,
                     Index col_lo = 0, Index col_hi = -1)
        : mtx_file_(mtx_file), idx_file_(idx_file), B_(batch_size)
    {
        mmutil::index::mtx_header_t h;
        ASSERT(mmutil::index::peek_bgzf_header(mtx_file, h) == EXIT_SUCCESS, "failed to read the header of " << mtx_file);
        D_ = h.max_row;   // rows = features, columns = data points (mmvae_io.hh:264-267)
        N_ = h.max_col;
        lo_ = std::max<Index>(0, col_lo);
        hi_ = col_hi < 0 ? N_ : std::min<Index>(N_, col_hi);
        ASSERT(lo_ < hi_, "empty shard [" << col_lo << ", " << col_hi << ") of " << N_ << " data points");
        ASSERT(mmutil::index::read_mmutil_index(idx_file, index_tab_) == EXIT_SUCCESS, "failed to read " << idx_file);
        if (const char *e = std::getenv("MMV_CSR_CACHE")) csr_cache = csr_cache || std::atoi(e) != 0;
        csr_cache = csr_cache && lo_ == 0 && hi_ == N_;   // the cache file describes the whole matrix
        if (csr_cache && load_cache()) {
            TLOG("Read " << idx_file_ << ".csr: " << D_ << " features x " << N_ << " data points, " << indices.size()
                         << " nonzeros (binary CSR cache of " << mtx_file << ")");
            return;
        }
        ASSERT(parse_all() == EXIT_SUCCESS, "failed to parse " << mtx_file);
        if (csr_cache && !save_cache()) WLOG("could not write " << idx_file_ << ".csr");
        TLOG("Read " << mtx_file << ": " << D_ << " features x " << N_ << " data points, " << indices.size()
                     << " nonzeros resident as CSR (" << ingest_threads_ << " reader threads: parse " << parse_seconds_
                     << " s, merge " << merge_seconds_ << " s)"
                     << (nlocal() == N_ ? std::string() : ", shard [" + std::to_string(lo_) + ", " + std::to_string(hi_) + ")"));
    }

    Index ntot() const { return N_; }
    Index nlocal() const { return hi_ - lo_; }     // data points held by this block (= ntot() unless it is a shard)
    Index shard_lo() const { return lo_; }
    Index size() const { return B_; }
    Index nfeature() const { return D_; }

    // A batch is a list of cell ids; nothing is copied (the reference fills a B x D buffer here, :208-245).
    void read(const std::vector<Index> &subcol) { batch_.assign(subcol.begin(), subcol.end()); }
    void clear() { batch_.clear(); }
    const std::vector<int64_t> &batch() const { return batch_; }

    // B x D dense view of a batch through the reference's own route: index table -> consecutive
    // blocks -> visit_bgzf_block -> last write wins, duplicates replicated (mmvae_io.hh:208-245).
    int dense_batch_indexed(const std::vector<Index> &subcol, std::vector<float> &mem) const
    {
        const size_t B = subcol.size();
        mem.assign(B * (size_t)D_, 0.f);
        std::vector<Index> sorted(subcol);
        std::sort(sorted.begin(), sorted.end());
        std::unordered_map<Index, Index> col2idx;
        for (Index cc : sorted)
            if (!col2idx.count(cc)) { const Index k = (Index)col2idx.size(); col2idx[cc] = k; }
        std::vector<std::vector<Index>> dup(B);
        for (size_t j = 0; j < B; ++j) dup[(size_t)col2idx[subcol[j]]].push_back((Index)j);
        for (const memory_block_t &blk : find_consecutive_blocks(index_tab_, subcol)) {
            const Index D = D_;
            int rc = mmutil::index::visit_bgzf_block(mtx_file_, blk.lb_mem, blk.ub_mem, [&](Index r, Index c, float w) {
                auto it = col2idx.find(c);
                if (it == col2idx.end() || r < 0 || r >= D) return;
                for (Index j : dup[(size_t)it->second]) mem[(size_t)j * (size_t)D + (size_t)r] = w;
            });
            if (rc != EXIT_SUCCESS) return rc;
        }
        return EXIT_SUCCESS;
    }

    // The same view from the resident CSR (what the device kernels see).
    void dense_batch_csr(const std::vector<Index> &subcol, std::vector<float> &mem) const
    {
        mem.assign(subcol.size() * (size_t)D_, 0.f);
        for (size_t j = 0; j < subcol.size(); ++j) {
            const Index c = subcol[j] - lo_;
            if (c < 0 || c >= nlocal()) continue;
            for (int64_t q = indptr[(size_t)c]; q < indptr[(size_t)c + 1]; ++q)
                mem[j * (size_t)D_ + (size_t)indices[(size_t)q]] = values[(size_t)q];
        }
    }

    // All data points as a dense [N x D] row-major matrix (covariate files: D = C is small).
    void dense_all(std::vector<float> &mem) const
    {
        std::vector<Index> all((size_t)nlocal());
        for (Index i = 0; i < nlocal(); ++i) all[(size_t)i] = lo_ + i;
        dense_batch_csr(all, mem);
    }

    const std::vector<Index> &index_tab() const { return index_tab_; }

    // Benchmark hook (--synthetic N,D,density; not in the reference): the shard [col_lo, col_hi) of an N x D count matrix
    // generated in memory instead of read from a file — cfg-4-sized runs of the mains need no 20 GB Matrix Market text.
    // Per cell the gene ids advance by 1 + a geometric gap (so they are increasing and distinct, ~density * D nonzeros),
    // counts are 1 + a small geometric; cell c's row depends only on (seed, c), so any sharding sees the same matrix.
    struct synthetic_t { };
    mtx_data_block_t(synthetic_t, Index ntot, Index nfeature, double density, const Index batch_size, uint64_t seed, Index col_lo = 0,
                     Index col_hi = -1)
        : mtx_file_("(synthetic)"), B_(batch_size), D_(nfeature), N_(ntot)
    {
        lo_ = std::max<Index>(0, col_lo);
        hi_ = col_hi < 0 ? N_ : std::min<Index>(N_, col_hi);
        ASSERT(lo_ < hi_ && D_ > 0 && density > 0. && density < 1., "--synthetic: empty shard or bad density");
        const Index n = hi_ - lo_;
        const int nthreads = (int)std::max<Index>(1, std::min<Index>(n / 1024, (Index)std::max(1u, std::thread::hardware_concurrency())));
        std::vector<std::vector<int32_t>> idx((size_t)nthreads);
        std::vector<std::vector<float>> val((size_t)nthreads);
        std::vector<std::vector<int32_t>> len((size_t)nthreads);
        const double lg = std::log1p(-density);
        auto work = [&](int t) {
            const Index c0 = lo_ + n * t / nthreads, c1 = lo_ + n * (t + 1) / nthreads;
            idx[(size_t)t].reserve((size_t)((double)(c1 - c0) * (double)D_ * density * 1.05) + 64);
            val[(size_t)t].reserve(idx[(size_t)t].capacity());
            for (Index c = c0; c < c1; ++c) {
                uint64_t st = (seed + 0x9E3779B97F4A7C15ull * (uint64_t)(c + 1)) | 1ull;
                auto next = [&]() {                                 // xorshift64*
                    st ^= st >> 12; st ^= st << 25; st ^= st >> 27;
                    return (double)((st * 2685821657736338717ull) >> 11) * (1.0 / 9007199254740992.0);
                };
                int32_t cnt = 0;
                for (double j = -1.;;) {
                    j += 1. + std::floor(std::log(1. - next()) / lg);   // geometric gap, P(gap = k) = (1 - d)^k d
                    if (j >= (double)D_) break;
                    idx[(size_t)t].push_back((int32_t)j);
                    val[(size_t)t].push_back(1.f + std::floor((float)(-1.4 * std::log(1. - next()))));
                    ++cnt;
                }
                len[(size_t)t].push_back(cnt);
            }
        };
        std::vector<std::thread> th;
        for (int t = 0; t < nthreads; ++t) th.emplace_back(work, t);
        for (auto &x : th) x.join();
        indptr.assign((size_t)n + 1, 0);
        size_t row = 0, total = 0;
        for (int t = 0; t < nthreads; ++t)
            for (int32_t l : len[(size_t)t]) { total += (size_t)l; indptr[++row] = (int64_t)total; }
        indices.reserve(total);
        values.reserve(total);
        for (int t = 0; t < nthreads; ++t) {
            indices.insert(indices.end(), idx[(size_t)t].begin(), idx[(size_t)t].end());
            values.insert(values.end(), val[(size_t)t].begin(), val[(size_t)t].end());
            std::vector<int32_t>().swap(idx[(size_t)t]);
            std::vector<float>().swap(val[(size_t)t]);
        }
        TLOG("Synthetic counts: " << D_ << " features x " << N_ << " data points at density " << density << ", shard [" << lo_ << ", "
                                  << hi_ << ") = " << total << " nonzeros resident as CSR");
    }

    std::vector<int64_t> indptr;    // [N + 1]
    std::vector<int32_t> indices;   // gene ids, increasing inside a cell
    std::vector<float> values;

    bool from_cache() const { return from_cache_; }

private:
    // ---- binary CSR cache: header {magic, D, N, nnz, mtx size, mtx mtime (ns), index-table checksum}, then
    //      indptr int64[N + 1], indices int32[nnz], values fp32[nnz].  Valid only for the byte-identical mtx file.
    struct cache_header_t {
        char magic[8];
        int64_t D, N, nnz, mtx_size, mtx_mtime_ns;
        uint64_t index_sum;
    };
    bool cache_key(cache_header_t &k) const
    {
        struct stat st;
        if (::stat(mtx_file_.c_str(), &st) != 0) return false;
        std::memset(&k, 0, sizeof(k));
        std::memcpy(k.magic, "MMVCSR1", 8);
        k.D = (int64_t)D_; k.N = (int64_t)N_; k.nnz = 0;
        k.mtx_size = (int64_t)st.st_size;
        k.mtx_mtime_ns = (int64_t)st.st_mtim.tv_sec * 1000000000LL + (int64_t)st.st_mtim.tv_nsec;
        uint64_t hsh = 1469598103934665603ull;                       // FNV-1a over the index table
        for (Index v : index_tab_) { hsh ^= (uint64_t)v; hsh *= 1099511628211ull; }
        k.index_sum = hsh;
        return true;
    }
    bool load_cache()
    {
        cache_header_t want, got;
        if (!cache_key(want)) return false;
        FILE *f = std::fopen((idx_file_ + ".csr").c_str(), "rb");
        if (!f) return false;
        bool ok = std::fread(&got, sizeof(got), 1, f) == 1 && std::memcmp(got.magic, want.magic, 8) == 0 && got.D == want.D &&
            got.N == want.N && got.mtx_size == want.mtx_size && got.mtx_mtime_ns == want.mtx_mtime_ns &&
            got.index_sum == want.index_sum && got.nnz >= 0;
        if (ok) {
            indptr.resize((size_t)N_ + 1); indices.resize((size_t)got.nnz); values.resize((size_t)got.nnz);
            ok = std::fread(indptr.data(), 8, indptr.size(), f) == indptr.size() &&
                std::fread(indices.data(), 4, indices.size(), f) == indices.size() &&
                std::fread(values.data(), 4, values.size(), f) == values.size() && std::fgetc(f) == EOF;
            ok = ok && indptr.front() == 0 && indptr.back() == got.nnz;
            for (size_t c = 0; ok && c < (size_t)N_; ++c) ok = indptr[c] <= indptr[c + 1];
            for (size_t q = 0; ok && q < indices.size(); ++q) ok = indices[q] >= 0 && indices[q] < (int32_t)D_;
        }
        std::fclose(f);
        if (!ok) { indptr.clear(); indices.clear(); values.clear(); }
        from_cache_ = ok;
        return ok;
    }
    bool save_cache() const
    {
        cache_header_t k;
        if (!cache_key(k)) return false;
        k.nnz = (int64_t)indices.size();
        const std::string path = idx_file_ + ".csr", tmp = path + ".tmp";
        FILE *f = std::fopen(tmp.c_str(), "wb");
        if (!f) return false;
        bool ok = std::fwrite(&k, sizeof(k), 1, f) == 1 && std::fwrite(indptr.data(), 8, indptr.size(), f) == indptr.size() &&
            std::fwrite(indices.data(), 4, indices.size(), f) == indices.size() &&
            std::fwrite(values.data(), 4, values.size(), f) == values.size();
        ok = (std::fclose(f) == 0) && ok;
        if (ok) ok = std::rename(tmp.c_str(), path.c_str()) == 0;   // readers never see a half-written cache
        if (!ok) std::remove(tmp.c_str());
        return ok;
    }
    bool from_cache_ = false;

    // One pass over the file into CSR.  The text is split at column boundaries taken from the index table
    // (bgzf virtual offsets, mmutil_index.hh:138-228) into one range per host thread; every thread inflates
    // and parses its own range with its own reader (mmutil_bgzf_util.hh:53-151 semantics), and the triplets
    // are merged in FILE order, so duplicates resolve exactly as in a sequential read (the last one wins,
    // mmvae_io.hh:127-130).  MMV_INGEST_THREADS overrides the thread count (1 = the sequential reader).
    int parse_all()
    {
        const Index N = N_, D = D_, lo = lo_, hi = hi_, NL = hi_ - lo_;
        // first valid (non-MISSING_POS) offset at or after a column: where that column's — or, when it has no
        // entries, the next present column's — text begins
        auto offset_at = [&](Index c) -> int64_t {
            while (c < N && c < (Index)index_tab_.size() && (int64_t)index_tab_[(size_t)c] == (int64_t)MISSING_POS) ++c;
            return (c < N && c < (Index)index_tab_.size()) ? (int64_t)index_tab_[(size_t)c] : 0;
        };
        // start right after the header line (not at voffset 0: the header itself has three fields)
        int64_t start = 0;
        if (lo == 0) {
            bgzf_reader_t fp;
            ERR_RET(!fp.open(mtx_file_), "unable to open " << mtx_file_);
            std::string line, tok[3];
            while (true) {
                ERR_RET(fp.getline(line) < 0, "Failed to read the header");
                if (line.empty() || line[0] == '%') continue;
                if (split_triplet_line(line, tok) == 3) break;
            }
            start = fp.tell();
        } else {
            start = offset_at(lo);
            if (start == 0) {   // no entries at or after the shard's first column: an all-empty shard
                indptr.assign((size_t)NL + 1, 0);
                indices.clear(); values.clear();
                return EXIT_SUCCESS;
            }
        }
        const int64_t stop = hi < N ? offset_at(hi) : 0;   // 0 = read to EOF
        int T = (int)std::thread::hardware_concurrency();
        if (const char *e = std::getenv("MMV_INGEST_THREADS")) T = std::atoi(e);
        T = std::max(1, std::min(T, 64));
        // range boundaries: strictly increasing, valid (non-zero) offsets of evenly spaced columns
        std::vector<int64_t> cut{ start };
        for (int t = 1; t < T; ++t) {
            Index c = lo + (Index)((int64_t)NL * t / T);
            while (c < hi && (int64_t)index_tab_[(size_t)c] <= cut.back()) ++c;   // skips MISSING_POS = 0 and gap-filled repeats
            if (c >= hi) break;
            if (stop != 0 && (int64_t)index_tab_[(size_t)c] >= stop) break;
            cut.push_back((int64_t)index_tab_[(size_t)c]);
        }
        const int R = (int)cut.size();
        const auto t_begin = std::chrono::steady_clock::now();
        // ---- fast path: a Matrix Market file indexed by column has its triplets in non-decreasing column order (the
        // reference's index builder insists on it, mmutil_index.hh), and the ranges are cut at column starts, so every
        // thread can append straight into a local CSR piece — gene ids and values of one data point after the other,
        // each tidied (sorted by gene, last duplicate wins) the moment the next one begins.  8 bytes per nonzero in the
        // pieces + 8 in the final arrays, no global counting sort.  A range that meets a column going backwards raises
        // `disorder` and the general path below redoes the work (24 bytes per nonzero at its peak).
        struct piece_t {
            std::vector<int32_t> g;
            std::vector<float> w;
            std::vector<std::pair<Index, int64_t>> cells;   // (shard-local data point, entries kept), ascending
            Index cur = -1;
            size_t cur_begin = 0;
            bool disorder = false;
            void finish()
            {
                if (cur < 0) return;
                const size_t n = g.size() - cur_begin;
                bool inc = true;
                for (size_t i = 1; i < n && inc; ++i) inc = g[cur_begin + i - 1] < g[cur_begin + i];
                size_t k = n;
                if (!inc) {
                    struct ent_t { int32_t g; float w; };
                    std::vector<ent_t> v(n);
                    for (size_t i = 0; i < n; ++i) v[i] = ent_t{ g[cur_begin + i], w[cur_begin + i] };
                    std::stable_sort(v.begin(), v.end(), [](const ent_t &a, const ent_t &b) { return a.g < b.g; });
                    k = 0;
                    for (size_t i = 0; i < n; ++i) {
                        if (i + 1 < n && v[i + 1].g == v[i].g) continue;
                        g[cur_begin + k] = v[i].g; w[cur_begin + k] = v[i].w; ++k;
                    }
                    g.resize(cur_begin + k); w.resize(cur_begin + k);
                }
                if (k > 0) cells.emplace_back(cur, (int64_t)k);
            }
        };
        {
            std::vector<piece_t> piece((size_t)R);
            std::vector<int> rc((size_t)R, EXIT_SUCCESS), bad((size_t)R, 0);
            auto parse_piece = [&](int t) {
                piece_t &pc = piece[(size_t)t];
                int &b = bad[(size_t)t];
                rc[(size_t)t] = mmutil::index::visit_bgzf_block(mtx_file_, cut[(size_t)t], t + 1 < R ? cut[(size_t)t + 1] : stop,
                                                                [&pc, &b, D, N, lo, hi](Index r, Index c, float w) {
                    if (r < 0 || r >= D || c < 0 || c >= N) { b = 1; return; }
                    if (c < lo || c >= hi || pc.disorder) return;
                    const Index cl = c - lo;
                    if (cl != pc.cur) {
                        if (cl < pc.cur) { pc.disorder = true; return; }
                        pc.finish();
                        pc.cur = cl;
                        pc.cur_begin = pc.g.size();
                    }
                    pc.g.push_back((int32_t)r);
                    pc.w.push_back(w);
                });
                if (!pc.disorder) pc.finish();
            };
            if (R == 1) parse_piece(0);
            else {
                std::vector<std::thread> th;
                for (int t = 0; t < R; ++t) th.emplace_back(parse_piece, t);
                for (auto &x : th) x.join();
            }
            bool ordered = true;
            for (int t = 0; t < R; ++t) {
                if (rc[(size_t)t] != EXIT_SUCCESS) return rc[(size_t)t];
                ERR_RET(bad[(size_t)t], "matrix market entry outside the " << D_ << " x " << N_ << " header shape in " << mtx_file_);
                ordered = ordered && !piece[(size_t)t].disorder;
                // (ranges are cut at column starts: a piece never reaches back into its predecessor's data points)
                if (t > 0 && !piece[(size_t)t].cells.empty() && !piece[(size_t)t - 1].cells.empty())
                    ordered = ordered && piece[(size_t)t - 1].cells.back().first < piece[(size_t)t].cells.front().first;
            }
            if (ordered) {
                const auto t_parsed = std::chrono::steady_clock::now();
                indptr.assign((size_t)NL + 1, 0);
                std::vector<int64_t> base((size_t)R + 1, 0);
                for (int t = 0; t < R; ++t) {
                    for (auto &ck : piece[(size_t)t].cells) indptr[(size_t)ck.first + 1] = ck.second;
                    base[(size_t)t + 1] = base[(size_t)t] + (int64_t)piece[(size_t)t].g.size();
                }
                for (Index c = 0; c < NL; ++c) indptr[(size_t)c + 1] += indptr[(size_t)c];
                indices.resize((size_t)base[(size_t)R]);
                values.resize((size_t)base[(size_t)R]);
                auto place = [&](int t) {
                    piece_t &pc = piece[(size_t)t];
                    std::copy(pc.g.begin(), pc.g.end(), indices.begin() + base[(size_t)t]);
                    std::copy(pc.w.begin(), pc.w.end(), values.begin() + base[(size_t)t]);
                    std::vector<int32_t>().swap(pc.g);
                    std::vector<float>().swap(pc.w);
                };
                if (R == 1) place(0);
                else {
                    std::vector<std::thread> th;
                    for (int t = 0; t < R; ++t) th.emplace_back(place, t);
                    for (auto &x : th) x.join();
                }
                ingest_threads_ = R;
                const auto t_end = std::chrono::steady_clock::now();
                parse_seconds_ = std::chrono::duration<double>(t_parsed - t_begin).count();
                merge_seconds_ = std::chrono::duration<double>(t_end - t_parsed).count();
                return EXIT_SUCCESS;
            }
            WLOG(mtx_file_ << ": columns are not in non-decreasing order; re-reading through the general (sorting) path");
        }
        // ---- general path: any triplet order
        struct trip_t { Index c; int32_t g; float w; };
        std::vector<std::vector<trip_t>> part((size_t)R);
        std::vector<int> rc((size_t)R, EXIT_SUCCESS), bad((size_t)R, 0);
        auto parse_range = [&](int t) {
            std::vector<trip_t> &out = part[(size_t)t];
            int &b = bad[(size_t)t];
            rc[(size_t)t] = mmutil::index::visit_bgzf_block(mtx_file_, cut[(size_t)t], t + 1 < R ? cut[(size_t)t + 1] : stop,
                                                            [&out, &b, D, N, lo, hi](Index r, Index c, float w) {
                // the reference does not range-check rows (mmutil_bgzf_util.hh:139 -> out-of-bounds write); we fail
                if (r < 0 || r >= D || c < 0 || c >= N) { b = 1; return; }
                if (c < lo || c >= hi) return;                      // a line of a neighbouring shard (block granularity)
                out.push_back(trip_t{ c - lo, (int32_t)r, w });
            });
        };
        if (R == 1) parse_range(0);
        else {
            std::vector<std::thread> th;
            for (int t = 0; t < R; ++t) th.emplace_back(parse_range, t);
            for (auto &x : th) x.join();
        }
        for (int t = 0; t < R; ++t) {
            if (rc[(size_t)t] != EXIT_SUCCESS) return rc[(size_t)t];
            ERR_RET(bad[(size_t)t], "matrix market entry outside the " << D_ << " x " << N_ << " header shape in " << mtx_file_);
        }
        const auto t_parsed = std::chrono::steady_clock::now();
        // bucket by data point in file order (counting sort: stable)
        std::vector<int64_t> cnt((size_t)NL + 1, 0);
        for (auto &p : part)
            for (auto &e : p) ++cnt[(size_t)e.c + 1];
        for (Index c = 0; c < NL; ++c) cnt[(size_t)c + 1] += cnt[(size_t)c];
        struct ent_t { int32_t g; float w; };
        std::vector<ent_t> all((size_t)cnt[(size_t)NL]);
        {
            std::vector<int64_t> cur(cnt.begin(), cnt.end() - 1);
            for (auto &p : part) {
                for (auto &e : p) all[(size_t)cur[(size_t)e.c]++] = ent_t{ e.g, e.w };
                std::vector<trip_t>().swap(p);
            }
        }
        // per data point: sort by gene (stable), duplicates keep the last; column ranges in parallel
        std::vector<int64_t> kept((size_t)NL + 1, 0);
        auto tidy = [&](Index c0, Index c1) {
            for (Index c = c0; c < c1; ++c) {
                ent_t *v = all.data() + cnt[(size_t)c];
                const size_t n = (size_t)(cnt[(size_t)c + 1] - cnt[(size_t)c]);
                std::stable_sort(v, v + n, [](const ent_t &a, const ent_t &b) { return a.g < b.g; });
                size_t k = 0;
                for (size_t i = 0; i < n; ++i) {
                    if (i + 1 < n && v[i + 1].g == v[i].g) continue;
                    v[k++] = v[i];
                }
                kept[(size_t)c + 1] = (int64_t)k;
            }
        };
        auto for_ranges = [&](const std::function<void(Index, Index)> &f) {
            if (T == 1 || NL < 1024) { f(0, NL); return; }
            std::vector<std::thread> th;
            for (int t = 0; t < T; ++t) th.emplace_back(f, (Index)((int64_t)NL * t / T), (Index)((int64_t)NL * (t + 1) / T));
            for (auto &x : th) x.join();
        };
        for_ranges(tidy);
        indptr.assign((size_t)NL + 1, 0);
        for (Index c = 0; c < NL; ++c) indptr[(size_t)c + 1] = indptr[(size_t)c] + kept[(size_t)c + 1];
        indices.resize((size_t)indptr[(size_t)NL]);
        values.resize((size_t)indptr[(size_t)NL]);
        for_ranges([&](Index c0, Index c1) {
            for (Index c = c0; c < c1; ++c) {
                const ent_t *v = all.data() + cnt[(size_t)c];
                int64_t q = indptr[(size_t)c];
                for (int64_t i = 0; i < kept[(size_t)c + 1]; ++i, ++q) {
                    indices[(size_t)q] = v[i].g;
                    values[(size_t)q] = v[i].w;
                }
            }
        });
        ingest_threads_ = R;
        const auto t_end = std::chrono::steady_clock::now();
        parse_seconds_ = std::chrono::duration<double>(t_parsed - t_begin).count();
        merge_seconds_ = std::chrono::duration<double>(t_end - t_parsed).count();
        return EXIT_SUCCESS;
    }

public:
    int ingest_threads() const { return ingest_threads_; }
    double parse_seconds() const { return parse_seconds_; }   // inflate + text parse (parallel part)
    double merge_seconds() const { return merge_seconds_; }   // bucket + sort + CSR assembly

private:
    int ingest_threads_ = 1;
    double parse_seconds_ = 0., merge_seconds_ = 0.;
    std::string mtx_file_, idx_file_;
    Index B_ = 0, D_ = 0, N_ = 0, lo_ = 0, hi_ = 0;
    std::vector<Index> index_tab_;
    std::vector<int64_t> batch_;
};

// create_ones_like (mmvae_io.hh:292-310) + write_matrix_market_stream (io.hh:191-226): a 1 x ntot sparse
// matrix of ones as bgzf Matrix Market text, triplets "1 <col> 1" sorted by column (1-based).
inline int
create_ones_like(int64_t ntot, const std::string &out_file)
{
    std::string text = "%%MatrixMarket matrix coordinate integer general\n";
    text += "1 " + std::to_string(ntot) + " " + std::to_string(ntot) + "\n";
    text.reserve(text.size() + (size_t)ntot * 12);
    for (int64_t j = 0; j < ntot; ++j) {
        text += "1 ";
        text += std::to_string(j + 1);
        text += " 1\n";
    }
    ERR_RET(!bgzf_write_file(out_file, text), "unable to write " << out_file);
    return EXIT_SUCCESS;
}

} // namespace mmvae
#endif
