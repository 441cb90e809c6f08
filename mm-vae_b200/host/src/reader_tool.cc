// mmv_reader_tool — test driver for the host reader and option parsers (tests/test_host_reader.py).
// Prints plain text or writes raw little-endian arrays so the CPU tests can compare them bit for bit
// with the golden output of the reference reader (tests/golden/reader_golden.npz).  No GPU calls.
#include "mmvae_io.hh"
#include "mmvae_options.hh"

#include <cinttypes>

static void
write_raw(const std::string &path, const void *p, size_t bytes)
{
    FILE *f = std::fopen(path.c_str(), "wb");
    if (!f) { ELOG("cannot write " << path); std::exit(1); }
    std::fwrite(p, 1, bytes, f);
    std::fclose(f);
}

int
main(const int argc, const char *argv[])
{
    if (argc < 2) { std::cerr << "usage: mmv_reader_tool <header|scan|table|block|dense|csr|shard|floats|write|ones|opts> ..." << std::endl; return 2; }
    const std::string cmd = argv[1];
    using mmutil::index::Index;
    if (cmd == "header" && argc == 3) {
        mmutil::index::mtx_header_t h;
        CHK(mmutil::index::peek_bgzf_header(argv[2], h));
        std::printf("%ld %ld %ld\n", (long)h.max_row, (long)h.max_col, (long)h.max_elem);
        std::printf("is_bgzf %d\n", mmvae::bgzf_reader_t::is_bgzf(argv[2]) ? 1 : 0);
    } else if (cmd == "scan" && argc == 3) {          // index text as build_mmutil_index writes it
        std::vector<std::pair<Index, int64_t>> c2f;
        CHK(mmutil::index::scan_mmutil_index(argv[2], c2f));
        for (auto &p : c2f) std::printf("%ld %ld\n", (long)p.first, (long)p.second);
    } else if (cmd == "table" && argc == 4) {         // build the index file, read it back with the gap fill
        CHK(mmutil::index::build_mmutil_index(argv[2], argv[3]));
        std::vector<Index> tab;
        CHK(mmutil::index::read_mmutil_index(argv[3], tab));
        for (Index v : tab) std::printf("%ld\n", (long)v);
    } else if (cmd == "block" && argc == 5) {         // row col float-bits
        CHK(mmutil::index::visit_bgzf_block(argv[2], std::stoll(argv[3]), std::stoll(argv[4]), [](Index r, Index c, float w) {
            uint32_t bits;
            std::memcpy(&bits, &w, 4);
            std::printf("%ld %ld %" PRIu32 "\n", (long)r, (long)c, bits);
        }));
    } else if (cmd == "dense" && argc >= 6) {         // dense <mtx> <idx> <out.bin> c0 c1 ... : indexed view, then CSR view
        mmvae::mtx_data_block_t blk(argv[2], argv[3], argc - 5);
        std::vector<Index> cols;
        for (int i = 5; i < argc; ++i) cols.push_back(std::stol(argv[i]));
        std::vector<float> a, b;
        CHK(blk.dense_batch_indexed(cols, a));
        blk.dense_batch_csr(cols, b);
        a.insert(a.end(), b.begin(), b.end());
        write_raw(argv[4], a.data(), a.size() * 4);
        std::printf("%ld %ld %ld\n", (long)blk.nfeature(), (long)blk.ntot(), (long)blk.indices.size());
    } else if (cmd == "csr" && (argc == 5 || argc == 6)) {   // csr <mtx> <idx> <out prefix> [cache]: last line says where it came from
        mmvae::mtx_data_block_t blk(argv[2], argv[3], 1, argc == 6);
        write_raw(std::string(argv[4]) + ".indptr", blk.indptr.data(), blk.indptr.size() * 8);
        write_raw(std::string(argv[4]) + ".indices", blk.indices.data(), blk.indices.size() * 4);
        write_raw(std::string(argv[4]) + ".values", blk.values.data(), blk.values.size() * 4);
        std::printf("%ld %ld %ld\n", (long)blk.nfeature(), (long)blk.ntot(), (long)blk.indices.size());
        if (argc == 6) std::printf("%s\n", blk.from_cache() ? "cache" : "parsed");
    } else if (cmd == "shard" && argc == 7) {         // shard <mtx> <idx> <out prefix> <lo> <hi>: CSR of data points [lo, hi) only
        mmvae::mtx_data_block_t blk(argv[2], argv[3], 1, false, std::stol(argv[5]), std::stol(argv[6]));
        write_raw(std::string(argv[4]) + ".indptr", blk.indptr.data(), blk.indptr.size() * 8);
        write_raw(std::string(argv[4]) + ".indices", blk.indices.data(), blk.indices.size() * 4);
        write_raw(std::string(argv[4]) + ".values", blk.values.data(), blk.values.size() * 4);
        std::printf("%ld %ld %ld %ld\n", (long)blk.nfeature(), (long)blk.ntot(), (long)blk.nlocal(), (long)blk.indices.size());
    } else if (cmd == "floats") {                     // take_float of each argument -> bits
        for (int i = 2; i < argc; ++i) {
            const float w = mmvae::take_float(argv[i], std::strlen(argv[i]));
            uint32_t bits;
            std::memcpy(&bits, &w, 4);
            std::printf("%" PRIu32 "\n", bits);
        }
    } else if (cmd == "write" && argc == 4) {         // write <text file> <out.gz>: bgzf-compress a text file
        FILE *f = std::fopen(argv[2], "rb");
        ERR_RET(!f, "cannot read " << argv[2]);
        std::string data;
        char buf[1 << 16];
        size_t n;
        while ((n = std::fread(buf, 1, sizeof(buf), f)) > 0) data.append(buf, n);
        std::fclose(f);
        ERR_RET(!mmvae::bgzf_write_file(argv[3], data), "cannot write " << argv[3]);
    } else if (cmd == "ones" && argc == 4) {          // ones <N> <out.mtx.gz>: create_ones_like + its index
        CHK(mmvae::create_ones_like(std::stoll(argv[2]), argv[3]));
        CHK(mmutil::index::build_mmutil_index(argv[3], std::string(argv[3]) + ".index"));
    } else if (cmd == "opts") {                       // parsed option values, one "key value" per line
        mmvae_options_t mo;
        training_options_t to;
        mmvae::nb::nbvae_options_t no;
        mmvae::vmf::vmf_options_t vo;
        const int rc = parse_mmvae_options(argc - 1, argv + 1, mo);
        CHK(parse_training_options(argc - 1, argv + 1, to));
        CHK(mmvae::nb::parse_nbvae_options(argc - 1, argv + 1, no));
        CHK(mmvae::vmf::parse_vmf_options(argc - 1, argv + 1, vo));
        auto join = [](const std::vector<int64_t> &v) { std::string s; for (auto x : v) s += (s.empty() ? "" : ",") + std::to_string(x); return s; };
        std::cout << "rc " << rc << "\nmtx " << mo.mtx << "\nidx " << mo.idx << "\nout " << mo.out << "\ncovar_mtx " << mo.covar_mtx
                  << "\ncovar_idx " << mo.covar_idx << "\nbatch_size " << mo.batch_size << "\nkl_discount " << mo.kl_discount
                  << "\nkl_min " << mo.kl_min << "\nkl_max " << mo.kl_max << "\nlr " << to.lr << "\ngrad_clip " << to.grad_clip
                  << "\nnboot " << to.nboot << "\nmax_epoch " << to.max_epoch << "\nrecording " << to.recording << "\nseed " << to.seed
                  << "\nngpu " << to.ngpu << "\nbf16 " << to.bf16 << "\ntrain_all_layers " << to.train_all_layers << "\nmean_encoding "
                  << join(no.mean_encoding_layers) << "\nmean_decoding " << join(no.mean_decoding_layers) << "\nmean_latent "
                  << no.mean_latent << "\noverdispersion_encoding " << no.overdispersion_encoding << "\noverdispersion_latent "
                  << no.overdispersion_latent << "\nnb_relu " << no.do_relu << "\nencoding " << join(vo.encoding_layers)
                  << "\ndecoding " << join(vo.decoding_layers) << "\nlatent " << vo.latent << "\nkappa_min " << vo.kappa_min
                  << "\nkappa_max " << vo.kappa_max << "\nvmf_relu " << vo.do_relu << "\nvmf_latent " << vo.vmf_latent << "\ncsr_cache " << mo.csr_cache
                  << std::endl;
    } else {
        std::cerr << "bad command" << std::endl;
        return 2;
    }
    return EXIT_SUCCESS;
}
