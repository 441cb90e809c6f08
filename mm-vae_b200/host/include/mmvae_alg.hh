// train_vae_model — the reference's epoch / batch / bootstrap schedule (include/mmvae_alg.hh:200-338)
// driving the C ABI.  Per batch: one forward in TRAIN mode (z is sampled, :274-279) whose loss is
// reported, then nboot x (bootstrap resample with replacement :292-301, forward, loss, zero_grad,
// backward, clip_grad_norm_, Adam step :303-310).  The last batch wraps modulo N (:264-266).  What the
// reference does per batch on the host — re-parse text, densify, copy to the device, index_select —
// is gone: a batch is the list of cell ids, the bootstrap is an index composition inside mmv_step.
//
// --ngpu G (not in the reference, which is single device): ONE process, one host thread + one mmv_ctx per
// GPU.  The data points are cut into G contiguous shards [g N / G, (g + 1) N / G); every thread ingests and
// uploads only its own shard (mtx_data_block_t with a column range), joins an in-process NCCL communicator
// (mmv_comm_unique_id / mmv_comm_init) and runs the same schedule on its shard: a global minibatch is
// batch_size cells PER GPU, gradients are averaged by the all-reduce inside mmv_step, every replica applies
// the same Adam update.  Rank 0 reports the (globally averaged) losses and writes the files; the recorder's
// latent table is filled shard by shard.
#ifndef MMVAE_B200_ALG_HH_
#define MMVAE_B200_ALG_HH_

#include "mmvae_model.hh"

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <iomanip>
#include <mutex>
#include <random>
#include <thread>
#include <type_traits>

// KL anneal shared by nb_loss_t / vmf_loss_t (nb_vae_main.cc:14-37, vmf_vae_main.cc:13-36):
// rate = max(max_rate * exp(-time_discount * epoch), min_rate)
struct kl_anneal_t {
    kl_anneal_t() : time_discount(0.f), max_rate(1.f), min_rate(1.f) { }
    float operator()(const int64_t epoch) const
    {
        const float t = static_cast<float>(epoch);
        const float rate = max_rate * std::exp(-time_discount * t);
        return std::max(rate, min_rate);
    }
    float time_discount, max_rate, min_rate;
};
// (the reference's mains define their own nb_loss_t / vmf_loss_t functors; those work with the six-argument
// train_vae_model below, this one with the device-aware overload and run_vae_main)

// One rank's share of train_vae_model.  `dev` says which shard this is (rank / nranks / sync); with one GPU
// it is the reference's loop verbatim.  Returns the epoch losses (meaningful on rank 0).
template <typename MODEL, typename VISITOR, typename DATA_BLOCK, typename LOSS>
std::vector<float>
train_vae_model(MODEL &model, VISITOR &visitor, DATA_BLOCK &data_block, mmvae::device_t &dev, training_options_t &opt,
                LOSS loss_fun)
{
    const bool lead = dev.rank == 0;
    if (lead) TLOG("Training on GPU " << opt.device << (dev.nranks > 1 ? " .. " + std::to_string(opt.device + dev.nranks - 1) : std::string())
                                      << " (B200, sm_100a)");
    const int64_t ntot = data_block.nlocal();          // this rank's data points (= ntot() on one GPU)
    const int64_t batch_size = data_block.size();
    ASSERT(batch_size > 0 && ntot > 0, "empty data or batch");
    // every rank runs the same number of minibatches: that of the largest shard (shorter shards wrap, :264-266)
    const int64_t nmax = dev.nranks > 1 ? (data_block.ntot() + dev.nranks - 1) / dev.nranks : ntot;
    int64_t nbatch = nmax / batch_size;
    if ((nbatch * batch_size) < nmax) ++nbatch;
    std::vector<typename DATA_BLOCK::Index> batch(batch_size);
    if (lead) TLOG("Batch size = " << batch_size << (dev.nranks > 1 ? " per GPU" : "") << ", Number of batches = " << nbatch);

    ASSERT(model.configure(opt) == EXIT_SUCCESS, "model options");
    if (lead) model.pretty_print(std::cerr);

    std::random_device rd;
    std::mt19937 rng(opt.seed ? (unsigned)(opt.seed + (uint64_t)dev.rank) : rd());
    std::uniform_int_distribution<> rboot(0, (int)batch_size - 1);
    std::vector<int64_t> ridx(batch_size);

    // --eps_file: injected noise, consumed in forward-call order
    std::vector<float> eps_stream;
    size_t eps_pos = 0;
    const size_t eps_n = (size_t)batch_size * (size_t)model.eps_per_cell;
    if (!opt.eps_file.empty()) {
        ASSERT(dev.nranks == 1, "--eps_file is a single-GPU reproducibility hook");
        FILE *f = std::fopen(opt.eps_file.c_str(), "rb");
        ASSERT(f, "unable to read " << opt.eps_file);
        std::fseek(f, 0, SEEK_END);
        const long bytes = std::ftell(f);
        std::fseek(f, 0, SEEK_SET);
        eps_stream.resize((size_t)bytes / 4);
        ASSERT(std::fread(eps_stream.data(), 4, eps_stream.size(), f) == eps_stream.size(), "short read: " << opt.eps_file);
        std::fclose(f);
    }
    auto next_eps = [&]() -> const float * {
        if (eps_stream.empty()) return nullptr;
        ASSERT(eps_pos + eps_n <= eps_stream.size(), "--eps_file holds fewer draws than the schedule makes");
        const float *p = eps_stream.data() + eps_pos;
        eps_pos += eps_n;
        return p;
    };
    const int eps_flag = eps_stream.empty() ? MMV_EPS_PHILOX : 0;

    std::vector<float> loss_vec;
    loss_vec.reserve(opt.max_epoch);
    const auto t_begin = std::chrono::steady_clock::now();
    auto t_warm = t_begin;

    for (int64_t epoch = 0; epoch < opt.max_epoch; ++epoch) {
        const float kl_weight = loss_fun(epoch);
        float _loss_epoch = 0.;
        for (int64_t b = 0; b < nbatch; ++b) {
            const int64_t lb = b * batch_size, ub = (b + 1) * batch_size;
            for (int64_t j = 0; j < (ub - lb); ++j) batch[j] = (lb + j) % ntot;
            data_block.read(batch);
            const std::vector<int64_t> &cells = data_block.batch();

            {   // reported loss: forward in train mode.  Deferred: this call hands back the terms of the previous
                // batch's evaluation, so the host never waits for the device inside the epoch (the reference pays
                // one .item() round trip per batch, mmvae_alg.hh:281); the last batch is collected after the loop
                mmv_loss_terms terms;
                MMV_OK_EXIT(mmv_step(model.get(), dev.csr, cells.data(), batch_size, nullptr, next_eps(), kl_weight,
                                     MMV_FWD | eps_flag | MMV_TERMS_DEFERRED, &terms));
                if (b > 0) {
                    _loss_epoch += terms.loss * batch_size;
                    if (lead) std::cerr << "\r[" << std::setw(20) << b << "] " << std::setw(20) << terms.loss;
                }
            }
            for (int64_t boot = 0; boot < opt.nboot; ++boot) {
                for (size_t r = 0; r < ridx.size(); ++r) ridx[r] = rboot(rng);
                MMV_OK_EXIT(mmv_step(model.get(), dev.csr, cells.data(), batch_size, ridx.data(), next_eps(), kl_weight,
                                     MMV_FWD | MMV_BWD | MMV_OPTIM | eps_flag, nullptr));
            }
            if ((epoch + 1) % opt.recording == 0) visitor.update_on_batch(model, data_block, batch, dev);
            data_block.clear();
        }
        {
            mmv_loss_terms terms;
            MMV_OK_EXIT(mmv_step_collect(model.get(), &terms));
            _loss_epoch += terms.loss * batch_size;
        }
        if (lead) std::cerr << "\r";
        _loss_epoch /= batch_size * nbatch;
        if (lead) TLOG("[" << std::setw(20) << (epoch + 1) << "] " << std::setw(20) << _loss_epoch);
        loss_vec.emplace_back(_loss_epoch);
        if (epoch == 0) {   // first epoch = warm-up (workspace allocation, communicator set-up): steady-state rate counts from here
            MMV_OK_EXIT(mmv_ctx_sync(dev.ctx));
            dev.barrier();
            t_warm = std::chrono::steady_clock::now();
        }
        if ((epoch + 1) % opt.recording == 0) {
            dev.barrier();                                   // every shard's rows of the latent table are in
            if (lead) visitor.update_on_epoch(model, epoch);
            dev.barrier();
        }
    }
    MMV_OK_EXIT(mmv_ctx_sync(dev.ctx));
    dev.barrier();
    if (lead) {
        const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
        const double uniq = (double)opt.max_epoch * nbatch * batch_size * dev.nranks;
        TLOG("Done training: " << opt.max_epoch << " epochs x " << nbatch << " minibatches x " << batch_size << " cells x "
                               << dev.nranks << " GPU(s) in " << sec << " s = " << uniq / sec << " unique cells/s under the schedule ("
                               << uniq * (double)opt.nboot / sec << " cells/s through the optimiser steps, nboot = " << opt.nboot << ")");
        if (opt.max_epoch > 1) {
            const double sec1 = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_warm).count();
            const double uniq1 = (double)(opt.max_epoch - 1) * nbatch * batch_size * dev.nranks;
            TLOG("  after the first epoch (warm-up): " << uniq1 / sec1 << " unique cells/s under the schedule, "
                                                        << uniq1 * (double)opt.nboot / sec1 << " cells/s through the optimiser steps");
        }
    }
    return loss_vec;
}

// The part of the two mains that does not depend on the model family (nb_vae_main.cc:39-133,
// vmf_vae_main.cc:38-127): index, covariates, per-GPU shard ingest + upload, model, training, score file.
// `make_model(ctx, n_features, c_dim)` returns the model (a std::unique_ptr to nbvae_t / vmf_vae_t).
template <typename MAKE_MODEL>
int
run_vae_main(const mmvae_options_t &main_options, training_options_t &train_opt, bool is_vmf, MAKE_MODEL make_model)
{
    const bool synthetic = main_options.syn_cells > 0;
    int64_t N = main_options.syn_cells;
    if (!synthetic) {
        if (!file_exists(main_options.mtx)) return EXIT_FAILURE;
        if (!file_exists(main_options.idx)) CHK(mmutil::index::build_mmutil_index(main_options.mtx, main_options.idx));
        mmutil::index::mtx_header_t hdr;
        CHK(mmutil::index::peek_bgzf_header(main_options.mtx, hdr));
        N = hdr.max_col;
    }

    // covariates (small: C x N) are read whole; each rank uploads its rows
    std::vector<float> covar;
    int64_t c_dim = 1;
    bool has_covar = false;
    if (file_exists(main_options.covar_mtx)) {
        if (!file_exists(main_options.covar_idx))
            CHK(mmutil::index::build_mmutil_index(main_options.covar_mtx, main_options.covar_idx));
        mmvae::mtx_data_block_t covar_block(main_options.covar_mtx, main_options.covar_idx, main_options.batch_size);
        ERR_RET(covar_block.ntot() != N, "data and covar on the same set of data points");
        covar_block.dense_all(covar);
        c_dim = covar_block.nfeature();
        has_covar = true;
    } else if (main_options.out.size() > 0 && !synthetic) {
        // nb_vae_main.cc:67-73: the reference writes a 1 x N all-ones Matrix Market file next to the outputs,
        // indexes it and reads it back; the same two files are produced here (create_ones_like), while the
        // device path keeps the all-ones covariate implicit (no B x 1 tensor, K4's tuned kernel stays eligible)
        const std::string covar_mtx = main_options.out + ".covar.mtx.gz";
        CHK(mmvae::create_ones_like(N, covar_mtx));
        TLOG("No covariate file is given. So we use this: " << covar_mtx);
        CHK(mmutil::index::build_mmutil_index(covar_mtx, covar_mtx + ".index"));
    }

    int G = std::max(1, train_opt.ngpu);
    if (G > 1) {
        int ndev = 0;
        MMV_OK(mmv_device_count(&ndev));
        ERR_RET(train_opt.device + G > ndev, "--ngpu " << G << " from device " << train_opt.device << ": only " << ndev << " GPU(s) visible");
        ERR_RET(N < G, "fewer data points than GPUs");
    }
    mmvae::shared_run_t shared(G);
    if (G > 1) MMV_OK(mmv_comm_unique_id(shared.uid));
    std::vector<float> scores;
    std::atomic<int> failed{ 0 };

    auto worker = [&](int g) {
        mmvae::device_t dev;
        dev.rank = g; dev.nranks = G; dev.shared = &shared;
        const int64_t lo = (int64_t)N * g / G, hi = (int64_t)N * (g + 1) / G;
        // ingest threads are shared between the ranks' readers
        std::unique_ptr<mmvae::mtx_data_block_t> data_p(
            synthetic ? new mmvae::mtx_data_block_t(mmvae::mtx_data_block_t::synthetic_t(), N, main_options.syn_genes,
                                                    main_options.syn_density, main_options.batch_size, train_opt.seed + 1234, lo, hi)
                      : new mmvae::mtx_data_block_t(main_options.mtx, main_options.idx, main_options.batch_size, main_options.csr_cache, lo, hi));
        mmvae::mtx_data_block_t &data = *data_p;
        std::vector<float> cshard;
        if (has_covar) cshard.assign(covar.begin() + lo * c_dim, covar.begin() + hi * c_dim);
        if (dev.upload(train_opt.device + g, data, has_covar ? &cshard : nullptr, c_dim) != EXIT_SUCCESS) { failed = 1; std::exit(1); }
        if (G > 1) MMV_OK_EXIT(mmv_comm_init(dev.ctx, G, g, shared.uid));
        if (g == 0) TLOG("Constructing a model");
        auto model = make_model(dev.ctx, data.nfeature(), c_dim);
        if (!train_opt.init_dir.empty()) ASSERT(model->load_params(train_opt.init_dir) == EXIT_SUCCESS, "--init_dir");
        if (g == 0) TLOG("Training the model...");
        mmvae::latent_recorder_t recorder(main_options.out, train_opt.max_epoch, &shared, N, is_vmf, train_opt.record_params);
        kl_anneal_t loss;
        loss.min_rate = main_options.kl_min;
        loss.max_rate = main_options.kl_max;
        loss.time_discount = main_options.kl_discount;
        auto s = train_vae_model(*model, recorder, data, dev, train_opt, loss);
        if (g == 0) scores = s;
        model.reset();          // before the context goes
    };
    if (G == 1) worker(0);
    else {
        std::vector<std::thread> th;
        for (int g = 0; g < G; ++g) th.emplace_back(worker, g);
        for (auto &t : th) t.join();
    }
    if (failed) return EXIT_FAILURE;
    CHK(mmvae::write_vector_file(main_options.out + ".scores.gz", scores));
    TLOG("Done");
    return EXIT_SUCCESS;
}

// train_vae_model with the reference's own signature (include/mmvae_alg.hh:200-213): model, visitor, data block, covariate
// block, options, loss functor — what a translation unit written against the reference calls (its two mains compile
// unchanged against host/include/refcompat, tests/test_refcompat_cpu.py).  Differences that remain, all forced by the
// absence of host tensors (INTEGRATION.md has the table):
//   * the model arrives unbound (dimensions only) and is bound to GPU opt.device here;
//   * `loss_fun(x, y, epoch)` is called with placeholders once per epoch and only its KL weight is used (the ELBO, its
//     backward and the optimiser run inside mmv_step);
//   * a visitor that is one of the reference's recorders (recorder_spec_t) becomes a latent_recorder_t; any other
//     visitor is called as update_on_batch(model, data_block, batch, dev) / update_on_epoch(model, epoch);
//   * one GPU (the data block was read whole by the caller); the sharded --ngpu run is run_vae_main's.
template <typename MODEL, typename VISITOR, typename DATA_BLOCK, typename LOSS>
std::vector<float>
train_vae_model(MODEL &model, VISITOR &visitor, DATA_BLOCK &data_block, DATA_BLOCK &covar_block, training_options_t &opt,
                LOSS loss_fun)
{
    ASSERT(data_block.ntot() == covar_block.ntot(), "data and covar on the same set of data points");
    ASSERT(data_block.size() == covar_block.size(), "data and covar on the same batch size");
    ASSERT(opt.ngpu <= 1, "--ngpu needs the sharded ingest of run_vae_main (each GPU reads its own range of the file)");
    // covariates: dense [N x C]; the all-ones single column the mains write when no --covar is given stays implicit
    std::vector<float> covar;
    covar_block.dense_all(covar);
    const int64_t c_dim = covar_block.nfeature();
    bool ones = c_dim == 1;
    for (size_t i = 0; ones && i < covar.size(); ++i) ones = covar[i] == 1.f;

    mmvae::shared_run_t shared(1);
    mmvae::device_t dev;
    dev.shared = &shared;
    ASSERT(dev.upload(opt.device, data_block, ones ? nullptr : &covar, c_dim) == EXIT_SUCCESS, "upload to GPU " << opt.device);
    ASSERT(model->bind(dev.ctx, opt.seed + 7) == EXIT_SUCCESS, "model construction on GPU " << opt.device);
    if (!opt.init_dir.empty()) ASSERT(model->load_params(opt.init_dir) == EXIT_SUCCESS, "--init_dir");
    auto kl = [&](const int64_t epoch) { return loss_fun(mmvae::batch_ref_t(), mmvae::recon_ref_t(), epoch).kl_weight; };
    std::vector<float> scores;
    if constexpr (std::is_base_of<mmvae::recorder_spec_t, VISITOR>::value) {
        mmvae::latent_recorder_t rec(visitor.header, visitor.max_epoch, &shared, data_block.ntot(), visitor.is_vmf, opt.record_params);
        scores = train_vae_model(*model.operator->(), rec, data_block, dev, opt, kl);
    } else {
        scores = train_vae_model(*model.operator->(), visitor, data_block, dev, opt, kl);
    }
    model->release();           // the model goes before its context (dev's destructor)
    return scores;
}

#endif
