// heat_b200/csrc/host/heat_host.hpp -- host-side C++ classes behind the `cf_c` module.
//
// They keep the names, constructor arguments and ownership rules of the reference classes that
// cf_cpu/src/pybind/init_modules.cpp binds (CFConfig cf_config.hpp:12-35; datasets::ClickDataset
// dataset.cpp:10-19; models::MatrixFactorization model.cpp:10-14; behavior_aggregators::
// AggregatorWeights behavior_aggregators.cpp:19-25; train::Engine engine.hpp:26-52; test::test_out
// test_out.cpp:9-11) but hold no compute: everything goes to the GPU through the C ABI of
// include/heat_b200.h.  Like the reference they BORROW the caller's numpy buffers (no copy, no
// keep-alive); unlike it every failure is reported (std::runtime_error -> Python exception).
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <vector>
#include "heat_b200.h"

namespace heat {
namespace host {

typedef uint64_t idx_t;   // splatt/base.h:49-51
typedef float val_t;      // CMakeLists.txt:11 (-DSPLATT_VAL_TYPEWIDTH=32)

struct CFConfig {
    idx_t emb_dim, num_negs, num_users, num_items, train_size, neg_sampler, tile_size, refresh_interval, num_subepochs;
    val_t l2, clip_val, l_r;
    std::vector<idx_t> milestones;
    // GPU-path knobs (defaults keep every existing YAML valid); settable as attributes from Python
    idx_t seed = 2022, sample_offset = 0;
    uint32_t chunk = 0, deterministic = 0, update_mode = 0, warps_per_block = 0, max_blocks = 0, sync_average = 0,
             history_mode = 0, engine = 0, tile_sampler = 0, exchange_mode = 0, sync_grads = 0, w0_tail_mode = 0, eval_precision = 0;
    CFConfig(idx_t emb_dim, idx_t num_negs, idx_t num_users, idx_t num_items, idx_t train_size, idx_t neg_sampler,
             idx_t tile_size, idx_t refresh_interval, idx_t num_subepochs, val_t l2, val_t clip_val,
             const std::vector<idx_t> &milestones, val_t l_r);
};

struct Dataset {
    virtual ~Dataset() {}
    idx_t data_rows = 0, data_cols = 0;
    const idx_t *data = nullptr;              // [data_rows, 2] (user, item)
    idx_t his_rows = 0, his_cols = 0;
    const idx_t *historical_items = nullptr;  // [num_users, max_his]
    const idx_t *masks = nullptr;             // [num_users, 1]
    int max_his = 0;                          // set by the frontend after construction (datasets.py:79)
};
struct ClickDataset : Dataset {
    ClickDataset(idx_t num_clicks, idx_t click_dim, const idx_t *clicks, idx_t his_rows, idx_t his_cols,
                 const idx_t *his, idx_t masks_rows, idx_t masks_cols, const idx_t *masks);
};

struct Model {
    virtual ~Model() {}
    std::shared_ptr<CFConfig> cf_config;
    val_t *user_weights = nullptr, *item_weights = nullptr;   // borrowed, trained in place
};
struct MatrixFactorization : Model {
    MatrixFactorization(std::shared_ptr<CFConfig> cfg, val_t *user_weights, val_t *item_weights);
};

struct AggregatorWeights {
    idx_t emb_dim;
    val_t *weights0;                                          // [emb_dim, emb_dim] borrowed
    AggregatorWeights(idx_t emb_dim, val_t *w0) : emb_dim(emb_dim), weights0(w0) {}
};

struct TestOut {
    void print(const std::string &s);
};

class Engine {
  public:
    Engine(std::shared_ptr<Dataset> dataset, std::shared_ptr<AggregatorWeights> aggregator_weights,
           std::shared_ptr<Model> model, std::shared_ptr<CFConfig> cf_config);
    ~Engine();
    Engine(const Engine &) = delete;
    Engine &operator=(const Engine &) = delete;

    val_t train_one_epoch();                  // engine.cpp:67 (upstream loop :293-347); host tables updated in place
    void evaluate0(val_t *out);               // engine.cpp:388-400; out = [num_users, num_items]
    // additive API
    void sync_to_host();
    void set_host_sync(bool on) { host_sync_ = on; }
    void evaluate_topk(uint32_t k, const int64_t *train_indptr, const int32_t *train_items, int32_t *ids, val_t *scores);
    void evaluate_metrics(const int64_t *train_indptr, const int32_t *train_items, const int64_t *test_indptr,
                          const int32_t *test_items, double out[6], double sums[7]);
    void comm_init(int rank, int world, const uint8_t uid[128]);
    void set_sync_interval(uint64_t samples);
    void save_checkpoint(const std::string &path);   // {tables, grad rows, W0, epoch, LR} -> one file
    void load_checkpoint(const std::string &path);   // restores the device state and the bound host tables
    void allreduce(double *vals, int n);
    idx_t epoch() const { return epoch_; }       // epochs trained so far (restored by load_checkpoint)
    float last_kernel_ms();
    float last_epoch_ms();
    float last_eval_ms();
    idx_t num_users() const { return cfg_->num_users; }
    idx_t num_items() const { return cfg_->num_items; }
    static std::vector<uint8_t> unique_id();

  private:
    void check(int rc, const char *what);
    std::shared_ptr<Dataset> dataset_;
    std::shared_ptr<AggregatorWeights> aggr_;
    std::shared_ptr<Model> model_;
    std::shared_ptr<CFConfig> cfg_;
    heat_ctx *ctx_ = nullptr;
    bool host_sync_ = true;
    idx_t epoch_ = 0;
};

}  // namespace host
}  // namespace heat
