// Option structs and parsers with the reference's field names, defaults, long options and aliases:
//   mmvae_options_t / parse_mmvae_options         include/mmvae.hh:31-209
//   training_options_t / parse_training_options   include/mmvae_alg.hh:14-125
//   nbvae_options_t / parse_nbvae_options         include/models/nb.hh:53-194
//   vmf_options_t / parse_vmf_options             include/models/vmf.hh:54-186
// Each parser scans the whole argv independently and ignores options it does not know, like the
// reference (getopt_long with opterr = 0).  Additions are long-only and default to the reference's
// behaviour: --seed, --device, --dtype, --train_all_layers, --grad_clip (declared but never parsed
// by the reference, mmvae_alg.hh:91-118), --overdisp_latent (named in the usage text only).
#ifndef MMVAE_B200_OPTIONS_HH_
#define MMVAE_B200_OPTIONS_HH_

#include "mmvae_util.hh"

#include <cstring>
#include <getopt.h>

namespace mmvae_detail {

// getopt permutes argv: run it over a private copy, as the reference does
struct argv_copy_t {
    std::vector<char *> ptr;
    std::vector<std::string> store;
    argv_copy_t(int argc, const char *argv[]) : store(argv, argv + argc)
    {
        for (auto &s : store) ptr.push_back(&s[0]);
        ptr.push_back(nullptr);
    }
};

template <typename F>
inline void
scan_options(int argc, const char *argv[], const char *short_opts, const option *long_opts, F &&on_opt)
{
    argv_copy_t copy(argc, argv);
    optind = 1;
    opterr = 0;
    while (true) {
        const int opt = getopt_long(argc, copy.ptr.data(), short_opts, long_opts, nullptr);
        if (opt == -1) break;
        on_opt(opt, optarg);
    }
}

} // namespace mmvae_detail

struct mmvae_options_t {
    explicit mmvae_options_t() : batch_size(100), kl_discount(.1f), kl_min(1e-2f), kl_max(1.f), csr_cache(false) { }
    std::string mtx, idx, out, row, col, annot;
    std::string covar_mtx, covar_idx;
    int64_t batch_size;
    float kl_discount, kl_min, kl_max;
    bool csr_cache;   // NOT in the reference (long option only): keep the parsed CSR as ${idx}.csr for the next run
    // NOT in the reference (long option only, benchmark hook): --synthetic N,D,density trains on an N x D count matrix
    // generated in memory (mmvae_io.hh) instead of --mtx, so the mains run at cfg-4 size without a 20 GB text file
    int64_t syn_cells = 0, syn_genes = 0;
    double syn_density = 0.;
};

inline int
parse_mmvae_options(const int argc, const char *argv[], mmvae_options_t &options)
{
    const char *usage = "\n[options]\n\n"
                        "--mtx         : a matrix market mtx file\n"
                        "--idx         : an index file for the mtx (default: ${mtx}.index)\n"
                        "--row         : the rows (line = one string)\n"
                        "--col         : the columns (line = one string)\n"
                        "--annot       : the list of column annotations (line = ${col} <space> ${k})\n"
                        "--out         : output file header\n"
                        "--covar       : a separate matrix market mtx for other covariates\n"
                        "--covar_idx   : an index file for the covar mtx (default: ${covar}.index)\n"
                        "--batch_size  : #samples in each batch (default: 100)\n"
                        "--csr_cache   : keep the parsed counts as ${idx}.csr and reuse it while the mtx file is unchanged\n"
                        "--synthetic   : N,D,density - train on counts generated in memory instead of --mtx (benchmark hook)\n\n"
                        "--kl_discount : KL divergence discount (default: 0)\n"
                        "              : Loss = likelihood_loss + beta * KL_loss\n"
                        "              : where beta = exp(- ${discount} * epoch)\n"
                        "--kl_max      : max KL divergence penalty (default: 1)\n"
                        "--kl_min      : min KL divergence penalty (default: 1e-2)\n\n";
    const option long_opts[] = { { "mtx", required_argument, nullptr, 'M' },        { "idx", required_argument, nullptr, 'I' },
                                 { "out", required_argument, nullptr, 'O' },        { "output", required_argument, nullptr, 'O' },
                                 { "cov", required_argument, nullptr, 'V' },        { "covar", required_argument, nullptr, 'V' },
                                 { "cov_idx", required_argument, nullptr, 'J' },    { "covar_idx", required_argument, nullptr, 'J' },
                                 { "row", required_argument, nullptr, 'r' },        { "col", required_argument, nullptr, 'c' },
                                 { "column", required_argument, nullptr, 'c' },     { "annot", required_argument, nullptr, 'a' },
                                 { "annotation", required_argument, nullptr, 'a' }, { "batch_size", required_argument, nullptr, 'b' },
                                 { "batch", required_argument, nullptr, 'b' },      { "kl_discount", required_argument, nullptr, 'K' },
                                 { "kl_max", required_argument, nullptr, 'L' },     { "kl_min", required_argument, nullptr, 'l' },
                                 { "csr_cache", no_argument, nullptr, 1002 },       { "csr-cache", no_argument, nullptr, 1002 },
                                 { "synthetic", required_argument, nullptr, 1003 },
                                 { "help", no_argument, nullptr, 'h' },             { nullptr, no_argument, nullptr, 0 } };
    bool help = false;
    mmvae_detail::scan_options(argc, argv, "M:I:O:V:J:r:c:a:b:K:L:l:h?", long_opts, [&](int opt, const char *arg) {
        switch (opt) {
        case 'M': options.mtx = arg; break;
        case 'I': options.idx = arg; break;
        case 'V': options.covar_mtx = arg; break;
        case 'J': options.covar_idx = arg; break;
        case 'O': options.out = arg; break;
        case 'r': options.row = arg; break;
        case 'c': options.col = arg; break;
        case 'a': options.annot = arg; break;
        case 'b': options.batch_size = std::stol(arg); break;
        case 'K': options.kl_discount = std::stof(arg); break;
        case 'l': options.kl_min = std::stof(arg); break;
        case 'L': options.kl_max = std::stof(arg); break;
        case 1002: options.csr_cache = true; break;
        case 1003: {
            long long n = 0, d = 0;
            double rho = 0.;
            if (std::sscanf(arg, "%lld,%lld,%lf", &n, &d, &rho) == 3) { options.syn_cells = n; options.syn_genes = d; options.syn_density = rho; }
            break;
        }
        case 'h': help = true; break;
        default: break;
        }
    });
    if (help) {
        std::cerr << usage << std::endl;
        return EXIT_SUCCESS;
    }
    ERR_RET(options.syn_cells <= 0 && !file_exists(options.mtx), "missing mtx file");
    ERR_RET(options.out.size() == 0, "need output file header");
    if (options.idx.size() == 0) options.idx = options.mtx + ".index";
    if (options.covar_idx.size() == 0) options.covar_idx = options.covar_mtx + ".index";
    return EXIT_SUCCESS;
}

struct training_options_t {
    explicit training_options_t()
        : lr(1e-3f), grad_clip(1.f), nboot(3), max_epoch(101), recording(10), device(0), seed(0), bf16(false), f32x3(false),
          train_all_layers(false), record_params(true), ngpu(1)
    {
    }
    float lr;          // 1e-3
    float grad_clip;   // 1
    int64_t nboot;     // 3
    int64_t max_epoch; // 101
    int64_t recording; // 10
    int device;        // CUDA ordinal (the reference picks kCUDA when available, mmvae_alg.hh:16-17)
    uint64_t seed;             // 0 = std::random_device like the reference (mmvae_alg.hh:242-243)
    bool bf16;                 // --dtype bf16: tcgen05 GEMMs (1e-2 tolerance); default fp32 (1e-4)
    bool f32x3;                // --dtype f32x3: the fp32 tolerance ON the tensor cores (operands as bf16 pairs)
    bool train_all_layers;     // default off = the reference's frozen unregistered Sequentials
    bool record_params;        // default on = the reference's per-parameter text dumps at every recorded epoch (nb.hh:599-615)
    int ngpu;                  // --ngpu G: cells sharded over G GPUs of this box, one host thread + one context each (default 1)
    // reproducibility hooks (long options only; the reference has neither: it seeds from random_device and draws
    // eps inside forward).  --init_dir: ${dir}/${parameter key}.f32 raw float32 in the libtorch [rows x cols] layout
    // replaces that parameter's initial value; --eps_file: raw float32 noise consumed in forward-call order
    // (B x (mu latent [+ nu latent]) floats per forward) instead of the library's own Philox draws
    std::string init_dir, eps_file;
};

inline int
parse_training_options(const int argc, const char *argv[], training_options_t &options)
{
    const char *usage = "[Training algorithm options]\n\n"
                        "--lr         : learning rate (default: 1e-3)\n"
                        "--grad_clip  : gradient clip (default: 1)\n"
                        "--nboot      : #bootstrapped gradients (default: 3)\n"
                        "--max_epoch  : maximum #epoch (default: 101)\n"
                        "--recording  : recording interval (default: 10)\n"
                        "--seed, --device, --dtype {f32,bf16,f32x3}, --train_all_layers, --skip_param_record : B200 build additions\n"
                        "--ngpu       : shard the cells over this many GPUs of the box (default: 1)\n"
                        "--init_dir, --eps_file : reproducibility hooks (initial parameters / injected noise)\n\n";
    const option long_opts[] = { { "lr", required_argument, nullptr, 'L' },         { "learning", required_argument, nullptr, 'L' },
                                 { "learn_rate", required_argument, nullptr, 'L' }, { "learning_rate", required_argument, nullptr, 'L' },
                                 { "rate", required_argument, nullptr, 'L' },       { "grad_clip", required_argument, nullptr, 'G' },
                                 { "nboot", required_argument, nullptr, 'B' },      { "boot", required_argument, nullptr, 'B' },
                                 { "bootstrap", required_argument, nullptr, 'B' },  { "max_epoch", required_argument, nullptr, 'E' },
                                 { "epoch", required_argument, nullptr, 'E' },      { "recording", required_argument, nullptr, 'R' },
                                 { "seed", required_argument, nullptr, 1001 },      { "device", required_argument, nullptr, 1002 },
                                 { "dtype", required_argument, nullptr, 1003 },     { "train_all_layers", no_argument, nullptr, 1004 },
                                 { "skip_param_record", no_argument, nullptr, 1005 },
                                 { "ngpu", required_argument, nullptr, 1006 },      { "init_dir", required_argument, nullptr, 1007 },
                                 { "eps_file", required_argument, nullptr, 1008 },
                                 { "help", no_argument, nullptr, 'h' },             { nullptr, no_argument, nullptr, 0 } };
    bool help = false;
    mmvae_detail::scan_options(argc, argv, "L:G:B:E:R:h", long_opts, [&](int opt, const char *arg) {
        switch (opt) {
        case 'L': options.lr = std::stof(arg); break;
        case 'G': options.grad_clip = std::stof(arg); break;
        case 'B': options.nboot = std::stol(arg); break;
        case 'E': options.max_epoch = std::stol(arg); break;
        case 'R': options.recording = std::stol(arg); break;
        case 1001: options.seed = std::stoull(arg); break;
        case 1002: options.device = std::stoi(arg); break;
        case 1003: options.bf16 = std::strcmp(arg, "bf16") == 0; options.f32x3 = std::strcmp(arg, "f32x3") == 0; break;
        case 1004: options.train_all_layers = true; break;
        case 1005: options.record_params = false; break;
        case 1006: options.ngpu = std::max(1, std::stoi(arg)); break;
        case 1007: options.init_dir = arg; break;
        case 1008: options.eps_file = arg; break;
        case 'h': help = true; break;
        default: break;
        }
    });
    if (help) std::cerr << usage << std::endl;
    return EXIT_SUCCESS;
}

namespace mmvae {
namespace nb {

struct nbvae_options_t {
    explicit nbvae_options_t() : mean_latent(2), overdispersion_encoding(1), overdispersion_latent(1), do_relu(false) { }
    std::vector<int64_t> mean_encoding_layers, mean_decoding_layers;
    int64_t mean_latent, overdispersion_encoding, overdispersion_latent;
    bool do_relu;
};

inline int
parse_nbvae_options(const int argc, const char *argv[], nbvae_options_t &options)
{
    const char *usage = "[Negative Binomial VAE options]\n\n"
                        "--mean_encoding     : dims for mean encoding layers (e.g., 10,10)\n"
                        "--mean_decoding     : dims for mean decoding layers (e.g., 10,10)\n\n"
                        "--mean_latent       : latent z's dim for the mean\n\n"
                        "--overdisp_encoding : dim for overdispersion encoding (default: 1)\n\n"
                        "--overdisp_latent   : latent dim for the overdispersion (default: 1)\n"
                        "--no_relu           : remove ReLU between layers (default)\n"
                        "--relu              : add ReLU between layers\n\n";
    const option long_opts[] = { { "mean_encoding", required_argument, nullptr, 'E' },
                                 { "mean-encoding", required_argument, nullptr, 'E' },
                                 { "mean_decoding", required_argument, nullptr, 'D' },
                                 { "mean-decoding", required_argument, nullptr, 'D' },
                                 { "mean_latent", required_argument, nullptr, 'L' },
                                 { "mean-latent", required_argument, nullptr, 'L' },
                                 { "overdisp_encoding", required_argument, nullptr, 'e' },
                                 { "overdisp-encoding", required_argument, nullptr, 'e' },
                                 { "overdispersion_encoding", required_argument, nullptr, 'e' },
                                 { "overdispersion-encoding", required_argument, nullptr, 'e' },
                                 { "overdispersion_latent", required_argument, nullptr, 'l' },
                                 { "overdispersion-latent", required_argument, nullptr, 'l' },
                                 { "overdisp_latent", required_argument, nullptr, 'l' },
                                 { "relu", no_argument, nullptr, 'R' },
                                 { "no_relu", no_argument, nullptr, 'r' },
                                 { "no-relu", no_argument, nullptr, 'r' },
                                 { "help", no_argument, nullptr, 'h' },
                                 { nullptr, no_argument, nullptr, 0 } };
    bool help = false;
    mmvae_detail::scan_options(argc, argv, "E:D:L:e:l:rRh", long_opts, [&](int opt, const char *arg) {
        switch (opt) {
        case 'E': copy_int_arr(arg, options.mean_encoding_layers); break;
        case 'D': copy_int_arr(arg, options.mean_decoding_layers); break;
        case 'L': options.mean_latent = std::stol(arg); break;
        case 'e': options.overdispersion_encoding = std::stol(arg); break;
        case 'l': options.overdispersion_latent = std::stol(arg); break;
        case 'r': options.do_relu = false; break;
        case 'R': options.do_relu = true; break;
        case 'h': help = true; break;
        default: break;
        }
    });
    if (help) std::cerr << usage << std::endl;
    return EXIT_SUCCESS;
}

} // namespace nb

namespace vmf {

struct vmf_options_t {
    explicit vmf_options_t() : latent(2), kappa_min(.1f), kappa_max(10.f), do_relu(false), vmf_latent(false) { }
    std::vector<int64_t> encoding_layers, decoding_layers;
    int64_t latent;
    float kappa_min, kappa_max;
    bool do_relu;
    bool vmf_latent;   // NOT in the reference: true vMF latent (Wood sampler, Householder, Bessel KL), long option only
};

inline int
parse_vmf_options(const int argc, const char *argv[], vmf_options_t &options)
{
    const char *usage = "[von Mises-Fisher VAE options]\n\n"
                        "--encoding  : dims for encoding layers (e.g., 10,10)\n"
                        "--decoding  : dims for decoding layers (e.g., 10,10)\n"
                        "--latent    : latent z's dim\n"
                        "--kappa_min : minimum concentration (default: 0.1)\n"
                        "--kappa_max : maximum concentration (default: 10)\n"
                        "--no_relu   : remove ReLU between layers (default)\n"
                        "--relu      : add ReLU between layers\n"
                        "--vmf_latent: von Mises-Fisher latent instead of the Gaussian one (needs --latent >= 3;\n"
                        "              not in the reference: Wood sampler, Householder reflection, Bessel KL)\n\n";
    const option long_opts[] = { { "encoding", required_argument, nullptr, 'E' },  { "decoding", required_argument, nullptr, 'D' },
                                 { "vmf_latent", no_argument, nullptr, 1001 },     { "vmf-latent", no_argument, nullptr, 1001 },
                                 { "latent", required_argument, nullptr, 'L' },    { "kappa_min", required_argument, nullptr, 'k' },
                                 { "kappa-min", required_argument, nullptr, 'k' }, { "kappa_max", required_argument, nullptr, 'K' },
                                 { "kappa-max", required_argument, nullptr, 'K' }, { "relu", no_argument, nullptr, 'R' },
                                 { "no_relu", no_argument, nullptr, 'r' },         { "no-relu", no_argument, nullptr, 'r' },
                                 { "help", no_argument, nullptr, 'h' },            { nullptr, no_argument, nullptr, 0 } };
    bool help = false;
    mmvae_detail::scan_options(argc, argv, "E:D:L:k:K:rRh", long_opts, [&](int opt, const char *arg) {
        switch (opt) {
        case 'E': copy_int_arr(arg, options.encoding_layers); break;
        case 'D': copy_int_arr(arg, options.decoding_layers); break;
        case 'r': options.do_relu = false; break;
        case 'R': options.do_relu = true; break;
        case 'L': options.latent = std::stol(arg); break;
        case 'k': options.kappa_min = std::stof(arg); break;
        case 'K': options.kappa_max = std::stof(arg); break;
        case 1001: options.vmf_latent = true; break;
        case 'h': help = true; break;
        default: break;
        }
    });
    if (help) std::cerr << usage << std::endl;
    return EXIT_SUCCESS;
}

} // namespace vmf
} // namespace mmvae
#endif
