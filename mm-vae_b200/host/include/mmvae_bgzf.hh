// BGZF container + the text number parsers of the reference reader, re-implemented on zlib.
//   container: tabix bgzf (reference src/ext/tabix/bgzf.c:193-267 writer, :385-424 read_block,
//              :559-579 seek, :626-665 getline; include/ext/tabix/bgzf.h:124 virtual offsets)
//   parsers  : strbuf_t::take_float / take_uint64 (include/utils/strbuf.hh:72-165) — bit-exact,
//              including their float/double evaluation order
//   splitter : the whitespace state machine of visit_bgzf_block (mmutil_bgzf_util.hh:111-127)
#ifndef MMVAE_B200_BGZF_HH_
#define MMVAE_B200_BGZF_HH_

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include <zlib.h>

namespace mmvae {

constexpr int64_t MISSING_POS = 0;          // mmutil_bgzf_util.hh:17
constexpr int BGZF_BLOCK_SIZE = 0xff00;     // bgzf.h:35: payload bytes per block written

class bgzf_reader_t {
public:
    bgzf_reader_t() = default;
    ~bgzf_reader_t() { close(); }
    bgzf_reader_t(const bgzf_reader_t &) = delete;

    bool open(const std::string &path)
    {
        close();
        fp_ = std::fopen(path.c_str(), "rb");
        file_pos_ = block_address_ = 0;
        block_offset_ = block_length_ = 0;
        return fp_ != nullptr;
    }
    void close()
    {
        if (fp_) std::fclose(fp_);
        fp_ = nullptr;
        if (zs_live_) inflateEnd(&zs_);
        zs_live_ = false;
        os_pos_ = -1;
    }
    static bool is_bgzf(const std::string &path)   // bgzf.c:581-593
    {
        static const unsigned char magic[16] = { 0x1f, 0x8b, 0x08, 0x04, 0, 0, 0, 0, 0, 0xff, 0x06, 0, 0x42, 0x43, 0x02, 0 };
        FILE *f = std::fopen(path.c_str(), "rb");
        if (!f) return false;
        unsigned char b[16];
        const size_t n = std::fread(b, 1, 16, f);
        std::fclose(f);
        return n == 16 && std::memcmp(b, magic, 16) == 0;
    }
    void seek(int64_t voffset)                     // bgzf.c:559-579
    {
        file_pos_ = block_address_ = voffset >> 16;
        block_offset_ = (int)(voffset & 0xFFFF);
        block_length_ = 0;
    }
    int64_t tell() const { return (block_address_ << 16) | (int64_t)(block_offset_ & 0xFFFF); }   // bgzf.h:124

    // bgzf_getline with '\n': returns the line length, or < 0 at EOF / error with nothing read
    int getline(std::string &out)
    {
        out.clear();
        int state = 0;
        while (state == 0) {
            if (block_offset_ >= block_length_) {
                if (read_block() != 0) { state = -2; break; }
                if (block_length_ == 0) { state = -1; break; }
            }
            const char *beg = block_.data() + block_offset_;
            const char *nl = (const char *)std::memchr(beg, '\n', (size_t)(block_length_ - block_offset_));
            int l;
            if (nl) { state = 1; l = (int)(nl - block_.data()); } else l = block_length_;
            out.append(beg, (size_t)(l - block_offset_));
            block_offset_ = l + 1;
            if (block_offset_ >= block_length_) {
                block_address_ = file_pos_;
                block_offset_ = 0;
                block_length_ = 0;
            }
        }
        if (out.empty() && state < 0) return state;
        return (int)out.size();
    }

private:
    int read_block()                               // bgzf.c:385-424
    {
        unsigned char hdr[18];
        const int64_t addr = file_pos_;
        if (os_pos_ != addr && fseeko(fp_, (off_t)addr, SEEK_SET) != 0) return -1;   // sequential reads never seek
        os_pos_ = -1;
        const size_t n = std::fread(hdr, 1, 18, fp_);
        if (n == 0) { block_length_ = 0; return 0; }
        if (n != 18 || hdr[0] != 0x1f || hdr[1] != 0x8b || hdr[2] != 8 || hdr[3] != 4 || hdr[12] != 'B' || hdr[13] != 'C')
            return -1;
        const int bsize = (int)(hdr[16] | (hdr[17] << 8)) + 1;
        if (bsize < 26) return -1;                  // 18-byte header + 8-byte trailer: anything shorter is corrupt or foreign
        comp_.resize((size_t)bsize - 18);
        if (std::fread(comp_.data(), 1, comp_.size(), fp_) != comp_.size()) return -1;
        block_.resize(0x10000);
        // one inflate state for the life of the reader (inflateReset per block instead of init + end: no 40 KB
        // allocation per 64 KB of text)
        if (!zs_live_) {
            std::memset(&zs_, 0, sizeof(zs_));
            if (inflateInit2(&zs_, -15) != Z_OK) return -1;
            zs_live_ = true;
        } else if (inflateReset(&zs_) != Z_OK) return -1;
        z_stream &zs = zs_;
        zs.next_in = (Bytef *)comp_.data();
        zs.avail_in = (uInt)(comp_.size() - 8);
        zs.next_out = (Bytef *)block_.data();
        zs.avail_out = (uInt)block_.size();
        const int rc = inflate(&zs, Z_FINISH);
        if (rc != Z_STREAM_END) return -1;
        {   // trailer: CRC32, ISIZE (little endian); ISIZE must be what inflate produced (<= 64 KiB)
            const unsigned char *tr = (const unsigned char *)comp_.data() + comp_.size() - 8;
            const uint32_t isize = (uint32_t)tr[4] | ((uint32_t)tr[5] << 8) | ((uint32_t)tr[6] << 16) | ((uint32_t)tr[7] << 24);
            if (isize > 0x10000u || isize != (uint32_t)zs.total_out) return -1;
        }
        file_pos_ = addr + bsize;
        os_pos_ = file_pos_;
        if (block_length_ != 0) block_offset_ = 0;   // kept when this read follows a seek
        block_address_ = addr;
        block_length_ = (int)zs.total_out;
        return 0;
    }

    FILE *fp_ = nullptr;
    z_stream zs_;
    bool zs_live_ = false;
    int64_t os_pos_ = -1;                          // where the FILE's position is known to be (-1: unknown)
    int64_t file_pos_ = 0, block_address_ = 0;
    int block_offset_ = 0, block_length_ = 0;
    std::vector<char> block_, comp_;
};

// Writer: blocks of BGZF_BLOCK_SIZE payload bytes + the 28-byte EOF marker (bgzf.c:193-267,487-535).
inline bool
bgzf_write_file(const std::string &path, const std::string &data)
{
    static const unsigned char eof_block[28] = { 0x1f, 0x8b, 0x08, 0x04, 0, 0, 0, 0, 0, 0xff, 0x06, 0, 0x42, 0x43,
                                                 0x02, 0, 0x1b, 0, 0x03, 0, 0, 0, 0, 0, 0, 0, 0, 0 };
    FILE *f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    std::vector<unsigned char> buf(0x10000 + 1024);
    for (size_t off = 0; off < data.size(); off += BGZF_BLOCK_SIZE) {
        const size_t len = std::min((size_t)BGZF_BLOCK_SIZE, data.size() - off);
        z_stream zs;
        std::memset(&zs, 0, sizeof(zs));
        if (deflateInit2(&zs, Z_DEFAULT_COMPRESSION, Z_DEFLATED, -15, 8, Z_DEFAULT_STRATEGY) != Z_OK) { std::fclose(f); return false; }
        zs.next_in = (Bytef *)(data.data() + off);
        zs.avail_in = (uInt)len;
        zs.next_out = buf.data() + 18;
        zs.avail_out = (uInt)(buf.size() - 18 - 8);
        const int rc = deflate(&zs, Z_FINISH);
        deflateEnd(&zs);
        if (rc != Z_STREAM_END) { std::fclose(f); return false; }
        const uint32_t bsize = 18 + (uint32_t)zs.total_out + 8;
        const unsigned char hdr[18] = { 0x1f, 0x8b, 0x08, 0x04, 0, 0, 0, 0, 0, 0xff, 0x06, 0, 'B', 'C', 0x02, 0,
                                        (unsigned char)((bsize - 1) & 0xff), (unsigned char)((bsize - 1) >> 8) };
        std::memcpy(buf.data(), hdr, 18);
        const uint32_t crc = (uint32_t)crc32(crc32(0L, Z_NULL, 0), (const Bytef *)(data.data() + off), (uInt)len);
        const uint32_t isize = (uint32_t)len;
        unsigned char *tail = buf.data() + 18 + zs.total_out;
        for (int i = 0; i < 4; ++i) { tail[i] = (crc >> (8 * i)) & 0xff; tail[4 + i] = (isize >> (8 * i)) & 0xff; }
        std::fwrite(buf.data(), 1, bsize, f);
    }
    std::fwrite(eof_block, 1, 28, f);
    std::fclose(f);
    return true;
}

// ---- text -> numbers -----------------------------------------------------------------------
inline int64_t
take_uint64(const char *s, size_t n)               // strbuf.hh:133-165: optional '-', leading digits only
{
    size_t i = 0;
    bool neg = false;
    int64_t x = 0;
    if (n > 0 && s[0] == '-') { neg = true; i = 1; }
    for (; i < n && s[i] >= '0' && s[i] <= '9'; ++i) x = x * 10 + (s[i] - '0');
    return neg ? -x : x;
}

inline float
take_float(const char *s, size_t n)                // strbuf.hh:72-131, same float/double evaluation order
{
    size_t i = 0;
    float r = 0.f;
    bool neg = false;
    if (i < n && s[i] == '-') { neg = true; ++i; }
    for (; i < n && s[i] >= '0' && s[i] <= '9'; ++i) r = (float)((double)r * 10.0 + (double)(s[i] - '0'));
    if (i < n && s[i] == '.') {
        float f = 0.f;
        int k = 0;
        for (++i; i < n && s[i] >= '0' && s[i] <= '9'; ++i, ++k) f = (float)((double)f * 10.0 + (double)(s[i] - '0'));
        r = (float)((double)r + (double)f / std::pow(10.0, (double)k));
    }
    if (i < n && (s[i] == 'e' || s[i] == 'E')) {
        ++i;
        float e = 0.f;
        bool neg_e = false;
        if (i < n && s[i] == '-') { neg_e = true; ++i; } else if (i < n && s[i] == '+') ++i;
        for (; i < n && s[i] >= '0' && s[i] <= '9'; ++i) e = (float)((double)e * 10.0 + (double)(s[i] - '0'));
        if (neg_e) e = -e;
        r = r * powf(10.f, e);
    }
    return neg ? -r : r;
}

// Whitespace splitter of visit_bgzf_block (mmutil_bgzf_util.hh:111-127): number of fields seen and the
// first three tokens (a token may be empty, like the reference's state machine produces).
inline int
split_triplet_line(const std::string &line, std::string tok[3])
{
    auto is_space = [](char c) { return c == ' ' || (c >= 9 && c <= 13); };
    tok[0].clear(); tok[1].clear(); tok[2].clear();
    std::string buf;
    int num_cols = 0;
    bool word = false;
    for (char ch : line) {
        if (is_space(ch) && !buf.empty()) {
            if (num_cols < 3) tok[num_cols] = buf;
            buf.clear();
            word = false;
            ++num_cols;
        } else {
            if (!is_space(ch)) buf.push_back(ch);
            word = true;
        }
    }
    if (word) {
        if (num_cols < 3) tok[num_cols] = buf;
        ++num_cols;
    }
    return num_cols;
}

} // namespace mmvae
#endif
