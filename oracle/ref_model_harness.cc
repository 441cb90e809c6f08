// TEST INFRASTRUCTURE ONLY (oracle/). Never linked into or called by the product.
//
// Compiles the REFERENCE's own model code where it lies under /root/reference
// (include/models/nb.hh, include/models/vmf.hh, include/operators.hh,
// include/modules/angular.hh) against libtorch CPU and
//   dump : runs the step of mmvae_alg.hh:274-311 with a fixed seed and fixed
//          bootstrap indices, writing params / eps / outputs / loss terms /
//          grads / post-Adam params as .npy files (golden vectors);
//   bench: times that step on the host cores (cpu_baseline / --impl reference);
//   train: the epoch x batch x bootstrap schedule of mmvae_alg.hh:254-334 around the reference's own model,
//          loss, Adam and clip calls, made reproducible (the reference seeds its bootstrap mt19937 from
//          std::random_device and draws eps inside forward): mt19937(seed) + uniform_int_distribution<> as in
//          :242-244, torch::manual_seed before every forward with the drawn eps saved in call order.  Golden
//          trajectory for the C++ mains (tests/golden/traj_*.npz).
// The outputs go to oracle/_ref/ (binary) and, via tests/golden/gen_golden.py,
// to tests/golden/*.npz.  No reference source is copied: it is #included.
#include <torch/torch.h>
#include <torch/csrc/autograd/variable.h>
#include <torch/csrc/autograd/function.h>
#include <torch/csrc/autograd/VariableTypeUtils.h>
#include <torch/csrc/autograd/functions/utils.h>
#define TORCH_INC_H_

#include "ref_shim/model_shim.hh"

namespace mmvae {
inline void write_tensor(const std::string, torch::Tensor) { }
}
using mmvae::write_tensor;

#include <cstring>
#include "util.hh"
#include "std_util.hh"
#include "check.hh"
#include "math.hh"
#include "nb.hh"
#include "vmf.hh"

#include <chrono>
#include <cstdio>
#include <fstream>
#include <map>
#include <random>
#include <sys/stat.h>

using torch::Tensor;

// ---------------------------------------------------------------- npy I/O
static void
save_npy(const std::string &path, const Tensor &t_)
{
    Tensor t = t_.detach().to(torch::kCPU).contiguous();
    std::string descr;
    if (t.scalar_type() == torch::kFloat32) descr = "<f4";
    else if (t.scalar_type() == torch::kFloat64) descr = "<f8";
    else if (t.scalar_type() == torch::kInt64) descr = "<i8";
    else { std::fprintf(stderr, "save_npy: dtype\n"); std::exit(1); }
    std::string shape = "(";
    for (int64_t i = 0; i < t.dim(); ++i)
        shape += std::to_string(t.size(i)) + ",";
    shape += ")";
    std::string hdr = "{'descr': '" + descr +
        "', 'fortran_order': False, 'shape': " + shape + ", }";
    size_t total = 10 + hdr.size() + 1;
    size_t pad = (64 - total % 64) % 64;
    hdr += std::string(pad, ' ');
    hdr += "\n";
    std::ofstream f(path, std::ios::binary);
    const char magic[] = "\x93NUMPY\x01\x00";
    f.write(magic, 8);
    uint16_t hl = (uint16_t)hdr.size();
    f.write((const char *)&hl, 2);
    f.write(hdr.data(), hdr.size());
    f.write((const char *)t.data_ptr(), t.numel() * t.element_size());
}

static Tensor
load_bin(const std::string &path, std::vector<int64_t> shape, torch::Dtype dt)
{
    Tensor t = torch::empty(shape, torch::TensorOptions().dtype(dt));
    std::ifstream f(path, std::ios::binary);
    if (!f) { std::fprintf(stderr, "cannot open %s\n", path.c_str()); std::exit(1); }
    f.read((char *)t.data_ptr(), t.numel() * t.element_size());
    if ((size_t)f.gcount() != (size_t)(t.numel() * t.element_size())) {
        std::fprintf(stderr, "short read %s\n", path.c_str()); std::exit(1);
    }
    return t;
}

static std::vector<int64_t>
parse_dims(const std::string &s)
{
    std::vector<int64_t> out;
    if (s.empty() || s == "-") return out;
    for (auto &tok : split(s, ',')) out.push_back(std::stol(tok));
    return out;
}

struct args_t {
    std::map<std::string, std::string> kv;
    std::string get(const std::string &k, const std::string &d = "") const
    {
        auto it = kv.find(k);
        return it == kv.end() ? d : it->second;
    }
    int64_t geti(const std::string &k, int64_t d) const
    {
        auto it = kv.find(k);
        return it == kv.end() ? d : std::stol(it->second);
    }
    double getf(const std::string &k, double d) const
    {
        auto it = kv.find(k);
        return it == kv.end() ? d : std::stod(it->second);
    }
};

// ------------------------------------------------------- model adapters
// Thin adapters so one driver serves both reference models.
struct nb_adapter {
    mmvae::nb::nbvae_t model { nullptr };
    int64_t R, Rnu;

    void build(const args_t &a, int64_t D, int64_t C)
    {
        using namespace mmvae::nb;
        std::vector<mu_encoder_h_dim> henc;
        std::vector<mu_decoder_h_dim> hdec;
        for (auto d : parse_dims(a.get("enc"))) henc.emplace_back((int)d);
        for (auto d : parse_dims(a.get("dec"))) hdec.emplace_back((int)d);
        R = a.geti("R", 2);
        Rnu = a.geti("Rnu", 1);
        // nb_vae_main.cc:103-112
        model = nbvae_t(data_dim((int)D), covar_dim((int)C), henc, hdec,
                        mu_encoder_r_dim((int)R),
                        nu_encoder_h_dim((int)a.geti("H", 1)),
                        nu_encoder_r_dim((int)Rnu), a.geti("relu", 0) != 0);
    }

    std::vector<std::pair<std::string, Tensor>> all_named()
    {
        std::vector<std::pair<std::string, Tensor>> out;
        for (auto &p : model->mu_enc->named_parameters(true))
            out.emplace_back(p.key(), p.value());
        for (auto &p : model->mu_dec->named_parameters(true))
            out.emplace_back(p.key(), p.value());
        for (auto &p : model->named_parameters(true))
            out.emplace_back(p.key(), p.value());
        return out;
    }

    std::vector<Tensor> eps_like(int64_t B)
    {   // same order as nb.hh:483 then :492
        Tensor e1 = torch::randn_like(torch::zeros({ B, R }));
        Tensor e2 = torch::randn_like(torch::zeros({ B, Rnu }));
        return { e1, e2 };
    }

    // returns total loss; fills named outputs
    Tensor fwd_loss(Tensor x, Tensor c, float w,
                    std::map<std::string, Tensor> *out)
    {
        auto y = model->forward(x, c);
        Tensor total = mmvae::nb::loss(x, y, w);
        if (out) {
            (*out)["recon_mu"] = y.recon_mu;
            (*out)["recon_nu"] = y.recon_nu;
            (*out)["recon_depth"] = y.recon_depth;
            (*out)["mu_mean"] = y.mu_mean;
            (*out)["mu_lnvar"] = y.mu_lnvar;
            (*out)["nu_mean"] = y.nu_mean;
            (*out)["nu_lnvar"] = y.nu_lnvar;
            (*out)["nllik"] = mmvae::nb::nllik_loss(x, y).reshape({ 1 });
            (*out)["kl_mu"] = mmvae::nb::kl_loss(y.mu_mean, y.mu_lnvar).reshape({ 1 });
            (*out)["kl_nu"] = mmvae::nb::kl_loss(y.nu_mean, y.nu_lnvar).reshape({ 1 });
            (*out)["loss"] = total.reshape({ 1 });
        }
        return total;
    }

    std::pair<Tensor, Tensor> encode(Tensor x) { return model->encode_mu(x); }
};

struct vmf_adapter {
    mmvae::vmf::vmf_vae_t model { nullptr };
    int64_t R;

    void build(const args_t &a, int64_t D, int64_t C)
    {
        using namespace mmvae::vmf;
        std::vector<z_enc_dim> henc;
        std::vector<z_dec_dim> hdec;
        for (auto d : parse_dims(a.get("enc"))) henc.emplace_back((int)d);
        for (auto d : parse_dims(a.get("dec"))) hdec.emplace_back((int)d);
        R = a.geti("R", 2);
        // vmf_vae_main.cc:100-107
        model = vmf_vae_t(data_dim((int)D), covar_dim((int)C), z_repr_dim((int)R),
                          henc, hdec, kappa_min((float)a.getf("kappa_min", .1)),
                          kappa_max((float)a.getf("kappa_max", 10.)),
                          a.geti("relu", 0) != 0);
        if (a.kv.count("ln_kappa")) {
            torch::NoGradGuard g;
            model->ln_kappa.fill_((float)a.getf("ln_kappa", 0.));
        }
    }

    std::vector<std::pair<std::string, Tensor>> all_named()
    {
        std::vector<std::pair<std::string, Tensor>> out;
        for (auto &p : model->z_enc->named_parameters(true))
            out.emplace_back(p.key(), p.value());
        for (auto &p : model->z_dec->named_parameters(true))
            out.emplace_back(p.key(), p.value());
        for (auto &p : model->named_parameters(true))
            out.emplace_back(p.key(), p.value());
        return out;
    }

    std::vector<Tensor> eps_like(int64_t B)
    {
        return { torch::randn_like(torch::zeros({ B, R })) };
    }

    Tensor fwd_loss(Tensor x, Tensor c, float w,
                    std::map<std::string, Tensor> *out)
    {
        auto y = model->forward(x, c);
        Tensor total = mmvae::vmf::vmf_vae_loss(x, y, w);
        if (out) {
            (*out)["recon"] = y.recon;
            (*out)["mean"] = y.mean;
            (*out)["lnvar"] = y.lnvar;
            (*out)["kappa"] = y.kappa.reshape({ 1 });
            (*out)["kl"] = mmvae::vmf::kl_loss_normal(y.mean, y.lnvar).reshape({ 1 });
            (*out)["loss"] = total.reshape({ 1 });
        }
        return total;
    }

    std::pair<Tensor, Tensor> encode(Tensor x) { return model->encode(x); }
};

static std::string
fname(const std::string &dir, const std::string &tag, std::string name)
{
    for (auto &ch : name) if (ch == '/') ch = '_';
    return dir + "/" + tag + "__" + name + ".npy";
}

template <typename A>
static std::vector<Tensor>
optim_params(A &ad, bool train_all)
{
    if (!train_all) return ad.model->parameters(); // mmvae_alg.hh:235
    std::vector<Tensor> ps;
    for (auto &p : ad.all_named()) ps.push_back(p.second);
    return ps;
}

// ------------------------------------------------------------------ dump
template <typename A>
static int
run_dump(const args_t &a)
{
    const int64_t D = a.geti("D", 0), C = a.geti("C", 1), B0 = a.geti("B0", 0);
    const int64_t B = a.geti("B", B0), nsteps = a.geti("steps", 1);
    const int64_t seed = a.geti("seed", 7);
    const float w = (float)a.getf("kl", 1.0);
    const bool train_all = a.geti("train_all", 0) != 0;
    const std::string out = a.get("out");
    ::mkdir(out.c_str(), 0755);

    Tensor x = load_bin(a.get("x"), { B0, D }, torch::kFloat32);
    Tensor c = load_bin(a.get("c"), { B0, C }, torch::kFloat32);
    Tensor ridx = load_bin(a.get("ridx"), { nsteps, B }, torch::kInt64);

    torch::manual_seed(seed);
    A ad;
    ad.build(a, D, C);
    for (auto &p : ad.all_named()) save_npy(fname(out, "init__param", p.first), p.second);

    // recorder path: eval-mode encode without covariates (nb.hh:627-628, vmf.hh:517)
    {
        ad.model->train(false);
        torch::NoGradGuard g;
        auto e = ad.encode(x);
        save_npy(fname(out, "enc", "mean"), e.first);
        save_npy(fname(out, "enc", "lnvar"), e.second);
    }

    auto params = optim_params(ad, train_all);
    torch::optim::Adam adam(params,
                            torch::optim::AdamOptions(a.getf("lr", 1e-3)).weight_decay(1e-4)); // mmvae_alg.hh:234-236
    ad.model->train(true); // mmvae_alg.hh:274

    for (int64_t s = 0; s < nsteps; ++s) {
        const std::string tag = "s" + std::to_string(s);
        Tensor xb = torch::index_select(x, 0, ridx[s]); // mmvae_alg.hh:300-301
        Tensor cb = torch::index_select(c, 0, ridx[s]);

        torch::manual_seed(seed + 1000 + s);
        auto eps = ad.eps_like(B);
        for (size_t k = 0; k < eps.size(); ++k)
            save_npy(fname(out, tag, "eps" + std::to_string(k)), eps[k]);
        torch::manual_seed(seed + 1000 + s);

        std::map<std::string, Tensor> outs;
        Tensor loss = ad.fwd_loss(xb, cb, w, &outs); // mmvae_alg.hh:303-304
        for (auto &kv : outs) save_npy(fname(out, tag + "__out", kv.first), kv.second);

        adam.zero_grad();   // :306
        loss.backward();    // :307
        for (auto &p : ad.all_named())
            if (p.second.grad().defined())
                save_npy(fname(out, tag + "__grad", p.first), p.second.grad());
        double gn = torch::nn::utils::clip_grad_norm_(params, a.getf("grad_clip", 1.0)); // :308
        save_npy(fname(out, tag, "gradnorm"), torch::full({ 1 }, gn, torch::kFloat64));
        adam.step();        // :310
        for (auto &p : ad.all_named()) save_npy(fname(out, tag + "__param", p.first), p.second);
    }
    return 0;
}

// ----------------------------------------------------------------- bench
template <typename A>
static int
run_bench(const args_t &a)
{
    const int64_t D = a.geti("D", 2000), C = a.geti("C", 1), B = a.geti("B", 100);
    const int64_t steps = a.geti("steps", 10), warmup = a.geti("warmup", 3);
    const double dens = a.getf("density", 0.05);
    const bool train_all = a.geti("train_all", 0) != 0;
    torch::manual_seed(a.geti("seed", 7));
    A ad;
    ad.build(a, D, C);
    Tensor x = (torch::rand({ B, D }) < dens).to(torch::kFloat32) *
        (1. + torch::poisson(torch::full({ B, D }, 1.5)));
    Tensor c = torch::ones({ B, C });
    auto params = optim_params(ad, train_all);
    torch::optim::Adam adam(params, torch::optim::AdamOptions(1e-3).weight_decay(1e-4));
    ad.model->train(true);
    auto step = [&]() {
        Tensor loss = ad.fwd_loss(x, c, 1.0f, nullptr);
        adam.zero_grad();
        loss.backward();
        torch::nn::utils::clip_grad_norm_(params, 1.0);
        adam.step();
        return loss.template item<float>();
    };
    float last = 0;
    for (int64_t i = 0; i < warmup; ++i) last = step();
    auto t0 = std::chrono::steady_clock::now();
    for (int64_t i = 0; i < steps; ++i) last = step();
    auto t1 = std::chrono::steady_clock::now();
    double sec = std::chrono::duration<double>(t1 - t0).count();
    std::printf("{\"ms_per_step\": %.6f, \"cells_per_s\": %.3f, \"threads\": %d, "
                "\"B\": %ld, \"D\": %ld, \"steps\": %ld, \"last_loss\": %.6g}\n",
                1e3 * sec / steps, (double)B * steps / sec, at::get_num_threads(),
                (long)B, (long)D, (long)steps, (double)last);
    return 0;
}


// ----------------------------------------------------------------- train
static void
save_raw(const std::string &path, const Tensor &t_)
{
    Tensor t = t_.detach().to(torch::kCPU).contiguous();
    std::ofstream f(path, std::ios::binary);
    f.write((const char *)t.data_ptr(), t.numel() * t.element_size());
}

template <typename A>
static int
run_train(const args_t &a)
{
    const int64_t D = a.geti("D", 0), C = 1, N = a.geti("N", 0), B = a.geti("B", 0);
    const int64_t epochs = a.geti("epochs", 2), nboot = a.geti("nboot", 3), seed = a.geti("seed", 7);
    const double kl_min = a.getf("kl_min", 1e-2), kl_max = a.getf("kl_max", 1.), kl_discount = a.getf("kl_discount", .1);
    const bool train_all = a.geti("train_all", 0) != 0;
    const std::string out = a.get("out");
    ::mkdir(out.c_str(), 0755);
    ::mkdir((out + "/init").c_str(), 0755);

    Tensor x_all = load_bin(a.get("x"), { N, D }, torch::kFloat32);
    Tensor c_all = torch::ones({ N, C });   // create_ones_like (nb_vae_main.cc:67-73)
    torch::manual_seed(seed);
    A ad;
    ad.build(a, D, C);
    for (auto &p : ad.all_named()) {
        std::string name = p.first;
        save_raw(out + "/init/" + name + ".f32", p.second);
        save_npy(fname(out, "init__param", name), p.second);
    }
    auto params = optim_params(ad, train_all);
    torch::optim::Adam adam(params, torch::optim::AdamOptions(a.getf("lr", 1e-3)).weight_decay(1e-4));   // :234-236
    const double grad_clip = a.getf("grad_clip", 1.0);

    std::mt19937 rng((unsigned)seed);                                   // :242-244 (seeded, not random_device)
    std::uniform_int_distribution<> rboot(0, (int)B - 1);
    std::vector<int64_t> ridx(B);
    int64_t nbatch = N / B;
    if (nbatch * B < N) ++nbatch;
    std::vector<Tensor> eps_stream, ridx_stream;
    std::vector<float> batch_loss, scores;
    int64_t fwd_k = 0;
    auto forward = [&](Tensor x, Tensor c, float w) {
        torch::manual_seed(seed + 1000 + fwd_k);
        for (auto &e : ad.eps_like(B)) eps_stream.push_back(e.reshape({ -1 }));
        torch::manual_seed(seed + 1000 + fwd_k);
        ++fwd_k;
        return ad.fwd_loss(x, c, w, nullptr);
    };
    for (int64_t epoch = 0; epoch < epochs; ++epoch) {
        const float w = (float)std::max(kl_max * std::exp(-kl_discount * (double)epoch), kl_min);   // nb_vae_main.cc:14-37
        float loss_epoch = 0.f;
        for (int64_t b = 0; b < nbatch; ++b) {
            std::vector<int64_t> batch(B);
            for (int64_t j = 0; j < B; ++j) batch[j] = (b * B + j) % N;                            // :264-266
            Tensor bi = torch::from_blob(batch.data(), { B }, torch::kInt64).clone();
            Tensor x = torch::index_select(x_all, 0, bi), c = torch::index_select(c_all, 0, bi);
            ad.model->train(true);
            {
                Tensor loss = forward(x, c, w).sum();                                             // :276-279
                const float lb = loss.template item<float>();
                loss_epoch += lb * B;
                batch_loss.push_back(lb);
            }
            for (int64_t boot = 0; boot < nboot; ++boot) {
                for (auto &r : ridx) r = rboot(rng);                                              // :294-295
                Tensor ri = torch::from_blob(ridx.data(), { B }, torch::kInt64).clone();
                ridx_stream.push_back(ri);
                Tensor loss = forward(torch::index_select(x, 0, ri), torch::index_select(c, 0, ri), w);
                adam.zero_grad();
                loss.backward();
                torch::nn::utils::clip_grad_norm_(params, grad_clip);
                adam.step();
            }
            ad.model->train(false);
        }
        scores.push_back(loss_epoch / (float)(B * nbatch));                                       // :323
    }
    save_raw(out + "/eps.f32", torch::cat(eps_stream));
    save_npy(out + "/eps.npy", torch::cat(eps_stream));
    save_npy(out + "/ridx.npy", torch::stack(ridx_stream));
    save_npy(out + "/batch_loss.npy", torch::from_blob(batch_loss.data(), { (int64_t)batch_loss.size() }, torch::kFloat32).clone());
    save_npy(out + "/scores.npy", torch::from_blob(scores.data(), { (int64_t)scores.size() }, torch::kFloat32).clone());
    for (auto &p : ad.all_named()) save_npy(fname(out, "final__param", p.first), p.second);
    return 0;
}

int
main(int argc, char **argv)
{
    if (argc < 2) {
        std::fprintf(stderr, "usage: %s dump|bench|train --key value ...\n", argv[0]);
        return 1;
    }
    args_t a;
    for (int i = 2; i + 1 < argc; i += 2) {
        std::string k = argv[i];
        if (k.rfind("--", 0) == 0) k = k.substr(2);
        a.kv[k] = argv[i + 1];
    }
    if (a.kv.count("threads")) at::set_num_threads((int)a.geti("threads", 1));
    const std::string mode = argv[1], model = a.get("model", "nb");
    if (mode == "dump")
        return model == "nb" ? run_dump<nb_adapter>(a) : run_dump<vmf_adapter>(a);
    if (mode == "bench")
        return model == "nb" ? run_bench<nb_adapter>(a) : run_bench<vmf_adapter>(a);
    if (mode == "train")
        return model == "nb" ? run_train<nb_adapter>(a) : run_train<vmf_adapter>(a);
    return 1;
}
