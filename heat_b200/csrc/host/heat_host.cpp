// heat_b200/csrc/host/heat_host.cpp -- see heat_host.hpp.  Host logic only; the GPU is reached
// through the extern "C" entry points of include/heat_b200.h.
#include "heat_host.hpp"

#include <cstdlib>
#include <cstring>
#include <iostream>
#include <stdexcept>

namespace heat {
namespace host {

CFConfig::CFConfig(idx_t emb_dim, idx_t num_negs, idx_t num_users, idx_t num_items, idx_t train_size, idx_t neg_sampler,
                   idx_t tile_size, idx_t refresh_interval, idx_t num_subepochs, val_t l2, val_t clip_val,
                   const std::vector<idx_t> &milestones, val_t l_r)
    : emb_dim(emb_dim), num_negs(num_negs), num_users(num_users), num_items(num_items), train_size(train_size),
      neg_sampler(neg_sampler), tile_size(tile_size), refresh_interval(refresh_interval), num_subepochs(num_subepochs),
      l2(l2), clip_val(clip_val), l_r(l_r), milestones(milestones)
{
}

ClickDataset::ClickDataset(idx_t num_clicks, idx_t click_dim, const idx_t *clicks, idx_t his_rows_, idx_t his_cols_,
                           const idx_t *his, idx_t masks_rows, idx_t masks_cols, const idx_t *masks_)
{
    if (click_dim != 2) throw std::invalid_argument("click_dataset must have shape [N, 2] (user, item)");
    if (masks_rows != his_rows_ || masks_cols != 1) throw std::invalid_argument("masks must have shape [num_users, 1]");
    data_rows = num_clicks; data_cols = click_dim; data = clicks;
    his_rows = his_rows_; his_cols = his_cols_; historical_items = his;
    masks = masks_;
    max_his = (int)his_cols_;
}

MatrixFactorization::MatrixFactorization(std::shared_ptr<CFConfig> cfg, val_t *uw, val_t *iw)
{
    cf_config = cfg; user_weights = uw; item_weights = iw;
}

void TestOut::print(const std::string &s) { std::cout << s << std::endl; }

void Engine::check(int rc, const char *what)
{
    if (rc != 0) throw std::runtime_error(std::string(what) + ": " + heat_last_error());
}

static int pick_device()
{
    for (const char *name : {"HEAT_DEVICE", "LOCAL_RANK"}) {
        const char *v = std::getenv(name);
        if (v && *v) return std::atoi(v);
    }
    return 0;
}

Engine::Engine(std::shared_ptr<Dataset> dataset, std::shared_ptr<AggregatorWeights> aggregator_weights,
               std::shared_ptr<Model> model, std::shared_ptr<CFConfig> cf_config)
    : dataset_(dataset), aggr_(aggregator_weights), model_(model), cfg_(cf_config)
{
    if (!dataset_ || !aggr_ || !model_ || !cfg_) throw std::invalid_argument("Engine: null argument");
    if (aggr_->emb_dim != cfg_->emb_dim) throw std::invalid_argument("Engine: aggregator weights are not [emb_dim, emb_dim]");
    if (dataset_->his_rows != cfg_->num_users)
        throw std::invalid_argument("Engine: historical_items has " + std::to_string(dataset_->his_rows) +
                                    " rows, config.num_users is " + std::to_string(cfg_->num_users));
    heat_config hc;
    std::memset(&hc, 0, sizeof(hc));
    hc.emb_dim = cfg_->emb_dim; hc.num_negs = cfg_->num_negs; hc.num_users = cfg_->num_users; hc.num_items = cfg_->num_items;
    hc.train_size = dataset_->data_rows; hc.neg_sampler = cfg_->neg_sampler; hc.tile_size = cfg_->tile_size;
    hc.refresh_interval = cfg_->refresh_interval; hc.num_subepochs = cfg_->num_subepochs;
    hc.l2 = cfg_->l2; hc.clip_val = cfg_->clip_val; hc.l_r = cfg_->l_r;
    hc.milestones = cfg_->milestones.data(); hc.n_milestones = cfg_->milestones.size();
    hc.seed = cfg_->seed; hc.sample_offset = cfg_->sample_offset; hc.chunk = cfg_->chunk;
    hc.deterministic = cfg_->deterministic; hc.update_mode = cfg_->update_mode; hc.warps_per_block = cfg_->warps_per_block;
    hc.max_blocks = cfg_->max_blocks; hc.sync_average = cfg_->sync_average;
    hc.history_mode = cfg_->history_mode; hc.engine = cfg_->engine; hc.tile_sampler = cfg_->tile_sampler; hc.exchange_mode = cfg_->exchange_mode;
    hc.sync_grads = cfg_->sync_grads; hc.w0_tail_mode = cfg_->w0_tail_mode; hc.eval_precision = cfg_->eval_precision;
    check(heat_create(&ctx_, &hc, pick_device()), "heat_create");
    try {
        // the stored column count is the layout; datasets.py:79 sets max_his to the same number
        check(heat_set_dataset(ctx_, dataset_->data, dataset_->data_rows, dataset_->historical_items, dataset_->his_rows,
                               dataset_->his_cols, dataset_->masks), "heat_set_dataset");
        check(heat_bind_tables(ctx_, model_->user_weights, model_->item_weights, aggr_->weights0), "heat_bind_tables");
    } catch (...) {
        heat_destroy(ctx_);
        ctx_ = nullptr;
        throw;
    }
}

Engine::~Engine()
{
    if (ctx_) heat_destroy(ctx_);
}

val_t Engine::train_one_epoch()
{
    double loss = 0.0;
    check(heat_train_epoch(ctx_, &loss), "heat_train_epoch");
    if (host_sync_) check(heat_sync_to_host(ctx_), "heat_sync_to_host");   // the in-place numpy contract
    epoch_ += 1;
    return (val_t)loss;
}

void Engine::sync_to_host() { check(heat_sync_to_host(ctx_), "heat_sync_to_host"); }
void Engine::evaluate0(val_t *out) { check(heat_eval_dense(ctx_, out), "heat_eval_dense"); }
void Engine::evaluate_topk(uint32_t k, const int64_t *train_indptr, const int32_t *train_items, int32_t *ids, val_t *scores)
{
    check(heat_eval_topk(ctx_, k, train_indptr, train_items, ids, scores), "heat_eval_topk");
}
void Engine::evaluate_metrics(const int64_t *train_indptr, const int32_t *train_items, const int64_t *test_indptr,
                              const int32_t *test_items, double out[6], double sums[7])
{
    check(heat_eval_metrics(ctx_, train_indptr, train_items, test_indptr, test_items, out, sums), "heat_eval_metrics");
}
void Engine::comm_init(int rank, int world, const uint8_t uid[128]) { check(heat_comm_init(ctx_, rank, world, uid), "heat_comm_init"); }
void Engine::set_sync_interval(uint64_t samples) { check(heat_set_sync_interval(ctx_, samples), "heat_set_sync_interval"); }
void Engine::allreduce(double *vals, int n) { check(heat_comm_allreduce_f64(ctx_, vals, n), "heat_comm_allreduce_f64"); }
void Engine::save_checkpoint(const std::string &path) { check(heat_save_state(ctx_, path.c_str()), "heat_save_state"); }
void Engine::load_checkpoint(const std::string &path)
{
    check(heat_load_state(ctx_, path.c_str()), "heat_load_state");
    uint64_t e = 0;
    check(heat_get_epoch(ctx_, &e), "heat_get_epoch");
    epoch_ = e;
}
float Engine::last_kernel_ms() { float v = 0; check(heat_last_kernel_ms(ctx_, &v), "heat_last_kernel_ms"); return v; }
float Engine::last_epoch_ms() { float v = 0; check(heat_last_epoch_ms(ctx_, &v), "heat_last_epoch_ms"); return v; }
float Engine::last_eval_ms() { float v = 0; check(heat_last_eval_ms(ctx_, &v), "heat_last_eval_ms"); return v; }

std::vector<uint8_t> Engine::unique_id()
{
    std::vector<uint8_t> id(128);
    if (heat_comm_unique_id(id.data()) != 0) throw std::runtime_error(std::string("heat_comm_unique_id: ") + heat_last_error());
    return id;
}

}  // namespace host
}  // namespace heat
