// oracle/ref_driver.cpp -- TEST INFRASTRUCTURE (CPU oracle), not product code.
//
// Thin C-ABI harness around the UNMODIFIED reference objects, compiled in place
// from /root/reference/cf_cpu/src by oracle/Makefile into oracle/_ref/.
// Nothing of the reference is copied here: this file only *calls*
//   cf::modules::models::MatrixFactorization::forward_backward   (matrix_factorization.cpp:15-181)
//   cf::modules::behavior_aggregators::BehaviorAggregator        (behavior_aggregators.cpp:28-153)
//   cf::modules::optimizers::SGD / Optimizer::scheduler_*         (sgd.cpp:14-26, optimizer.cpp:24-38)
//   cf::modules::negative_samplers::UniformRandomNegativeSampler  (uniform_random_negative_sampler.cpp:10-36)
//   cf::modules::train::Engine::evaluate0                         (engine.cpp:388-400)
// and restates the upstream thread-pool epoch loop that survives only as a
// comment block in the fork (engine.cpp:293-347), because the fork's live MPI
// driver (engine.cpp:67-386) needs MPI and never terminates at one rank
// (engine.cpp:231).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference legs may load the library built from this file.
#include <omp.h>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#include "modules/cf_config.hpp"
#include "modules/cf_modules.hpp"
#include "modules/datasets/click_dataset.hpp"
#include "modules/models/matrix_factorization.hpp"
#include "modules/behavior_aggregators/behavior_aggregator.hpp"
#include "modules/negative_samplers/uniform_random_negative_sampler.hpp"
#include "modules/negative_samplers/random_tile_negative_sampler.hpp"
#include "modules/optimizers/sgd.hpp"
#include "modules/train/engine.hpp"

#include "philox_ref.h"

using namespace cf::modules;

namespace {

struct RefHandle {
    std::shared_ptr<CFConfig> cfg;
    std::shared_ptr<datasets::ClickDataset> data;
    std::shared_ptr<behavior_aggregators::AggregatorWeights> w0;
    std::shared_ptr<models::MatrixFactorization> model;
    std::shared_ptr<CFModules> mods;
    // single-step mode keeps one aggregator / thread buffer alive across calls
    std::unique_ptr<behavior_aggregators::BehaviorAggregator> step_aggr;
    std::unique_ptr<memory::ThreadBuffer> step_buf;
    uint64_t epoch = 0;
};

}  // namespace

extern "C" {

void *heat_ref_create(uint64_t emb_dim, uint64_t num_negs, uint64_t num_users, uint64_t num_items,
                      uint64_t train_size, uint64_t max_his, float l2, float clip_val, float l_r,
                      const uint64_t *milestones, uint64_t n_milestones,
                      uint64_t *clicks, uint64_t *his_items, uint64_t *masks,
                      float *user_w, float *item_w, float *w0)
{
    auto *h = new RefHandle();
    std::vector<idx_t> ms(milestones, milestones + n_milestones);
    // ctor argument order: cf_config.hpp:14-17 (neg_sampler 0, tile 512, refresh 8192, subepochs 2)
    h->cfg = std::make_shared<CFConfig>(emb_dim, num_negs, num_users, num_items, train_size,
                                        (idx_t)0, (idx_t)512, (idx_t)8192, (idx_t)2,
                                        l2, clip_val, ms, l_r);
    h->data = std::make_shared<datasets::ClickDataset>(train_size, (idx_t)2, clicks, num_users, max_his,
                                                       his_items, num_users, (idx_t)1, masks);
    h->data->max_his = (int)max_his;  // datasets.py:79 sets it after construction
    h->w0 = std::make_shared<behavior_aggregators::AggregatorWeights>((int)emb_dim, w0);
    h->model = std::make_shared<models::MatrixFactorization>(h->cfg, user_w, item_w);
    h->mods = std::make_shared<CFModules>(h->cfg.get(), new optimizers::SGD(h->cfg));  // engine.cpp:16-17
    return h;
}

void heat_ref_destroy(void *hv) { delete static_cast<RefHandle *>(hv); }

void heat_ref_set_lr(void *hv, float l_r) { static_cast<RefHandle *>(hv)->mods->optimizer->l_r = l_r; }
float heat_ref_get_lr(void *hv) { return static_cast<RefHandle *>(hv)->mods->optimizer->l_r; }
uint64_t heat_ref_get_epoch(void *hv) { return static_cast<RefHandle *>(hv)->epoch; }

// One call of the reference's forward_backward with caller-supplied negatives.
// A single BehaviorAggregator instance persists across calls (like one OpenMP
// thread's instance inside an epoch); reset_aggregator starts a fresh one.
float heat_ref_step(void *hv, uint64_t user, uint64_t pos, const uint64_t *neg_ids, int reset_aggregator)
{
    auto *h = static_cast<RefHandle *>(hv);
    if (!h->step_aggr || reset_aggregator) {
        h->step_aggr.reset(new behavior_aggregators::BehaviorAggregator(h->data.get(), h->w0.get(), h->cfg));
        h->step_buf.reset(new memory::ThreadBuffer(h->cfg->emb_dim, h->cfg->num_negs));
    }
    std::vector<idx_t> negs(neg_ids, neg_ids + h->cfg->num_negs);
    return h->model->forward_backward(user, pos, negs, h->mods, h->step_buf.get(), h->step_aggr.get());
}

void heat_ref_get_grads(void *hv, float *gu, float *gi)
{
    auto *h = static_cast<RefHandle *>(hv);
    const size_t d = h->cfg->emb_dim;
    if (gu) std::memcpy(gu, h->model->user_embedding->grads->data, h->cfg->num_users * d * sizeof(float));
    if (gi) std::memcpy(gi, h->model->item_embedding->grads->data, h->cfg->num_items * d * sizeof(float));
}

void heat_ref_zero_grad(void *hv)
{
    auto *h = static_cast<RefHandle *>(hv);
    h->model->user_embedding->zero_grad();
    h->model->item_embedding->zero_grad();
}

// Restatement of the upstream epoch (engine.cpp:156-160 LR schedule, :294-342
// thread-pool loop, :345-347 zero_grad, :378-385 loss/N).
//   sampler 0: the reference's own UniformRandomNegativeSampler, seed (epoch+1)*tid (engine.cpp:302)
//   sampler 1: Philox contract of philox_ref.h fed through the reference's
//              ignore_pos logic with the persistent per-thread vector (nthreads must be 1
//              for this to be deterministic and equal to the stateless look-back rule)
//   sampler 2: Philox contract via the stateless look-back rule (any nthreads)
double heat_ref_train_epoch(void *hv, int sampler, uint64_t seed, uint64_t sample_offset, int nthreads)
{
    auto *h = static_cast<RefHandle *>(hv);
    const idx_t N = h->cfg->train_size;
    const idx_t K = h->cfg->num_negs;
    const idx_t I = h->cfg->num_items;
    if (h->cfg->milestones.size() > 1)
        h->mods->optimizer->scheduler_multi_step_lr(h->epoch, h->cfg->milestones, 0.1);
    else
        h->mods->optimizer->scheduler_step_lr(h->epoch, h->cfg->milestones[0], 0.1);

    if (nthreads <= 0) nthreads = omp_get_max_threads();
    double local_loss = 0.;
    const uint64_t *clicks = h->data->data->data;
    const uint64_t epoch = h->epoch;
    datasets::Dataset *train_data_ptr = h->data.get();
    behavior_aggregators::AggregatorWeights *w0_ptr = h->w0.get();

#pragma omp parallel num_threads(nthreads) reduction(+ : local_loss)
    {
        idx_t user_id = 0, pos_id = 0;
        std::vector<idx_t> neg_ids(K);
        const idx_t thread_id = omp_get_thread_num();
        const idx_t tseed = (epoch + 1) * thread_id;
        negative_samplers::UniformRandomNegativeSampler ref_sampler(h->cfg, tseed);
        memory::ThreadBuffer t_buf(h->cfg->emb_dim, K);
        behavior_aggregators::BehaviorAggregator aggr(train_data_ptr, w0_ptr, h->cfg);

#pragma omp for schedule(dynamic, 512)
        for (idx_t i = 0; i < N; ++i) {
            // click_dataset.cpp:17-22 uses a shared row_buf (racy under OpenMP); read directly.
            user_id = clicks[2 * i];
            pos_id = clicks[2 * i + 1];
            if (sampler == 0) {
                ref_sampler.ignore_pos_sampling(user_id, pos_id, neg_ids);
            } else if (sampler == 1) {
                for (idx_t k = 0; k < K; ++k) {
                    idx_t neg_id = heat_oracle_draw(seed, epoch, sample_offset + i, (uint32_t)k, I);
                    if (neg_id != pos_id) neg_ids[k] = neg_id;  // uniform_random_negative_sampler.cpp:31-34
                }
            } else {
                for (idx_t k = 0; k < K; ++k)
                    neg_ids[k] = heat_oracle_neg_id(seed, epoch, sample_offset, clicks, i, (uint32_t)k, I);
            }
            local_loss += h->model->forward_backward(user_id, pos_id, neg_ids, h->mods, &t_buf, &aggr);
        }
    }
    h->model->user_embedding->zero_grad();
    h->model->item_embedding->zero_grad();
    h->epoch += 1;
    return local_loss / (double)N;
}

// Reference evaluation: Engine::evaluate0 (engine.cpp:388-400), result copied out.
void heat_ref_evaluate0(void *hv, float *out)
{
    auto *h = static_cast<RefHandle *>(hv);
    train::Engine engine(h->data, h->w0, h->model, h->cfg);
    engine.evaluate0();
    std::memcpy(out, engine.sim_matrix.data(),
                (size_t)engine.sim_matrix.rows() * engine.sim_matrix.cols() * sizeof(float));
}

int heat_ref_max_threads(void) { return omp_get_max_threads(); }

}  // extern "C"
