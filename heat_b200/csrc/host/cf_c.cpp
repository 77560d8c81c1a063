// heat_b200/csrc/host/cf_c.cpp -- pybind11 module `cf_c`: the drop-in boundary.
//
// Same module tree, class names, keyword arguments and attributes as the reference's binding
// (cf_cpu/src/pybind/cf_c.cpp:6-9, init_modules.cpp:13-152), so the reference's Python frontend
// (cf_cpu/cf/*.py) runs against it unchanged:
//   cf_c.modules.CFConfig(emb_dim, num_negs, num_users, num_items, train_size, neg_sampler, tile_size,
//                         refresh_interval, num_subepoches, l2, clip_val, milestones, l_r)   .emb_dim
//   cf_c.modules.datasets.ClickDataset(click_dataset, historical_items, masks)   .data_rows .max_his
//   cf_c.modules.models.MatrixFactorization(cf_config, user_weights, item_weights)
//   cf_c.modules.behavior_aggregators.AggregatorWeights(aggregator_weights0)   .emb_dim
//   cf_c.modules.train.Engine(dataset, aggregator_weights, model, cf_config)
//        .train_one_epoch() -> float     .evaluate0() -> float32 [num_users, num_items]
//   cf_c.modules.test.test_out().test(str)
// Additive: Engine.evaluate_topk / evaluate_metrics / sync_to_host / set_host_sync / comm_init /
// set_sync_interval / allreduce / last_*_ms, train.unique_id(), extra CFConfig attributes.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include "heat_host.hpp"

namespace py = pybind11;
using namespace heat::host;

template <class T> using carray = py::array_t<T, py::array::c_style | py::array::forcecast>;

static void need_dims(const py::array &a, int nd, const char *name)
{
    if (a.ndim() != nd) throw std::invalid_argument(std::string(name) + " must be " + std::to_string(nd) + "-dimensional");
}

PYBIND11_MODULE(cf_c, cf_module)
{
    cf_module.doc() = "HEAT cf_c for B200 (heat_b200): same surface as CF_CPU pybind11";
    py::module_ m = cf_module.def_submodule("modules", "modules");

    py::class_<CFConfig, std::shared_ptr<CFConfig>>(m, "CFConfig")
        .def(py::init<idx_t, idx_t, idx_t, idx_t, idx_t, idx_t, idx_t, idx_t, idx_t, val_t, val_t, const std::vector<idx_t> &, val_t>(),
             py::arg("emb_dim"), py::arg("num_negs"), py::arg("num_users"), py::arg("num_items"), py::arg("train_size"),
             py::arg("neg_sampler"), py::arg("tile_size"), py::arg("refresh_interval"), py::arg("num_subepoches"),
             py::arg("l2"), py::arg("clip_val"), py::arg("milestones"), py::arg("l_r"))
        .def_readwrite("emb_dim", &CFConfig::emb_dim)
        .def_readwrite("seed", &CFConfig::seed)
        .def_readwrite("sample_offset", &CFConfig::sample_offset)
        .def_readwrite("chunk", &CFConfig::chunk)
        .def_readwrite("deterministic", &CFConfig::deterministic)
        .def_readwrite("update_mode", &CFConfig::update_mode)
        .def_readwrite("warps_per_block", &CFConfig::warps_per_block)
        .def_readwrite("max_blocks", &CFConfig::max_blocks)
        .def_readwrite("sync_average", &CFConfig::sync_average)
        .def_readwrite("history_mode", &CFConfig::history_mode)
        .def_readwrite("engine", &CFConfig::engine)
        .def_readwrite("tile_sampler", &CFConfig::tile_sampler)
        .def_readwrite("exchange_mode", &CFConfig::exchange_mode)
        .def_readwrite("sync_grads", &CFConfig::sync_grads)
        .def_readwrite("w0_tail_mode", &CFConfig::w0_tail_mode)
        .def_readwrite("eval_precision", &CFConfig::eval_precision);

    py::module_ datasets = m.def_submodule("datasets", "datasets");
    py::class_<Dataset, std::shared_ptr<Dataset>>(datasets, "Dataset");
    py::class_<ClickDataset, Dataset, std::shared_ptr<ClickDataset>>(datasets, "ClickDataset")
        .def(py::init([](py::array_t<idx_t, py::array::c_style> &click_dataset, py::array_t<idx_t, py::array::c_style> &historical_items,
                         py::array_t<idx_t, py::array::c_style> &masks) {
                 need_dims(click_dataset, 2, "click_dataset"); need_dims(historical_items, 2, "historical_items"); need_dims(masks, 2, "masks");
                 return std::make_shared<ClickDataset>(
                     (idx_t)click_dataset.shape(0), (idx_t)click_dataset.shape(1), static_cast<const idx_t *>(click_dataset.request().ptr),
                     (idx_t)historical_items.shape(0), (idx_t)historical_items.shape(1), static_cast<const idx_t *>(historical_items.request().ptr),
                     (idx_t)masks.shape(0), (idx_t)masks.shape(1), static_cast<const idx_t *>(masks.request().ptr));
             }),
             py::arg("click_dataset"), py::arg("historical_items"), py::arg("masks"))
        .def_readwrite("data_rows", &ClickDataset::data_rows)
        .def_readwrite("max_his", &ClickDataset::max_his);

    // additive: native LightGCN-format reader (heat_clickfile_*, include/heat_b200.h) behind datasets.ClickDataset
    datasets.def("parse_clicks", [](const std::string &path, const std::string &separator) {
        heat_clickfile *f = nullptr;
        uint64_t rows = 0, nnz = 0;
        if (heat_clickfile_open(path.c_str(), separator.empty() ? ' ' : separator[0], &f, &rows, &nnz) != 0)
            throw std::runtime_error(heat_last_error());
        py::array_t<int64_t> users((py::ssize_t)rows), indptr((py::ssize_t)rows + 1), items((py::ssize_t)nnz);
        const int rc = heat_clickfile_read(f, users.mutable_data(), indptr.mutable_data(), items.mutable_data());
        heat_clickfile_close(f);
        if (rc != 0) throw std::runtime_error(heat_last_error());
        return py::make_tuple(users, indptr, items);
    }, py::arg("path"), py::arg("separator") = " ");

    py::module_ models = m.def_submodule("models", "models");
    py::class_<Model, std::shared_ptr<Model>>(models, "Model");
    py::class_<MatrixFactorization, Model, std::shared_ptr<MatrixFactorization>>(models, "MatrixFactorization")
        .def(py::init([](const std::shared_ptr<CFConfig> cf_config, py::array_t<val_t, py::array::c_style> &user_weights,
                         py::array_t<val_t, py::array::c_style> &item_weights) {
                 need_dims(user_weights, 2, "user_weights"); need_dims(item_weights, 2, "item_weights");
                 if ((idx_t)user_weights.shape(0) != cf_config->num_users || (idx_t)item_weights.shape(0) != cf_config->num_items ||
                     (idx_t)user_weights.shape(1) != cf_config->emb_dim || (idx_t)item_weights.shape(1) != cf_config->emb_dim)
                     throw std::invalid_argument("user_weights / item_weights must be [num_users, emb_dim] / [num_items, emb_dim]");
                 return std::make_shared<MatrixFactorization>(cf_config, static_cast<val_t *>(user_weights.request().ptr),
                                                              static_cast<val_t *>(item_weights.request().ptr));
             }),
             py::arg("cf_config"), py::arg("user_weights"), py::arg("item_weights"));

    py::module_ aggr = m.def_submodule("behavior_aggregators", "Behavior aggregators");
    py::class_<AggregatorWeights, std::shared_ptr<AggregatorWeights>>(aggr, "AggregatorWeights")
        .def(py::init([](py::array_t<val_t, py::array::c_style> &aggregator_weights0) {
                 need_dims(aggregator_weights0, 2, "aggregator_weights0");
                 if (aggregator_weights0.shape(0) != aggregator_weights0.shape(1)) throw std::invalid_argument("aggregator_weights0 must be square");
                 return std::make_shared<AggregatorWeights>((idx_t)aggregator_weights0.shape(0), static_cast<val_t *>(aggregator_weights0.request().ptr));
             }),
             py::arg("aggregator_weights0"))
        .def_readwrite("emb_dim", &AggregatorWeights::emb_dim);

    py::module_ train = m.def_submodule("train", "train");
    train.def("unique_id", []() { auto id = Engine::unique_id(); return py::bytes(reinterpret_cast<const char *>(id.data()), id.size()); });
    py::class_<Engine, std::shared_ptr<Engine>>(train, "Engine")
        .def(py::init<std::shared_ptr<Dataset>, std::shared_ptr<AggregatorWeights>, std::shared_ptr<Model>, std::shared_ptr<CFConfig>>(),
             py::arg("dataset"), py::arg("aggregator_weights"), py::arg("model"), py::arg("cf_config"))
        .def("train_one_epoch", &Engine::train_one_epoch)
        .def("evaluate0", [](Engine &e) {
            py::array_t<val_t> out({(py::ssize_t)e.num_users(), (py::ssize_t)e.num_items()});
            e.evaluate0(static_cast<val_t *>(out.request().ptr));
            return out;
        })
        .def("sync_to_host", &Engine::sync_to_host)
        .def("set_host_sync", &Engine::set_host_sync, py::arg("on"))
        .def("evaluate_topk", [](Engine &e, uint32_t k, py::object train_indptr, py::object train_items) {
            py::array_t<int32_t> ids({(py::ssize_t)e.num_users(), (py::ssize_t)k});
            py::array_t<val_t> sc({(py::ssize_t)e.num_users(), (py::ssize_t)k});
            if (train_indptr.is_none()) {
                e.evaluate_topk(k, nullptr, nullptr, ids.mutable_data(), sc.mutable_data());
            } else {
                carray<int64_t> ip(train_indptr); carray<int32_t> it(train_items);
                if ((idx_t)ip.size() != e.num_users() + 1) throw std::invalid_argument("train_indptr must have num_users + 1 entries");
                e.evaluate_topk(k, ip.data(), it.data(), ids.mutable_data(), sc.mutable_data());
            }
            return py::make_tuple(ids, sc);
        }, py::arg("k"), py::arg("train_indptr") = py::none(), py::arg("train_items") = py::none())
        .def("evaluate_metrics", [](Engine &e, py::object train_indptr, py::object train_items, carray<int64_t> test_indptr, carray<int32_t> test_items) {
            if ((idx_t)test_indptr.size() != e.num_users() + 1) throw std::invalid_argument("test_indptr must have num_users + 1 entries");
            py::array_t<double> out(6), sums(7);
            if (train_indptr.is_none()) {
                e.evaluate_metrics(nullptr, nullptr, test_indptr.data(), test_items.data(), out.mutable_data(), sums.mutable_data());
            } else {
                carray<int64_t> ip(train_indptr); carray<int32_t> it(train_items);
                if ((idx_t)ip.size() != e.num_users() + 1) throw std::invalid_argument("train_indptr must have num_users + 1 entries");
                e.evaluate_metrics(ip.data(), it.data(), test_indptr.data(), test_items.data(), out.mutable_data(), sums.mutable_data());
            }
            return py::make_tuple(out, sums);
        }, py::arg("train_indptr"), py::arg("train_items"), py::arg("test_indptr"), py::arg("test_items"))
        .def("comm_init", [](Engine &e, int rank, int world, py::bytes uid) {
            std::string s = uid;
            if (s.size() != 128) throw std::invalid_argument("unique id must be 128 bytes");
            e.comm_init(rank, world, reinterpret_cast<const uint8_t *>(s.data()));
        }, py::arg("rank"), py::arg("world"), py::arg("unique_id"))
        .def("set_sync_interval", &Engine::set_sync_interval, py::arg("samples"))
        .def("allreduce", [](Engine &e, carray<double> vals) {
            py::array_t<double> out(vals.size());
            std::copy(vals.data(), vals.data() + vals.size(), out.mutable_data());
            e.allreduce(out.mutable_data(), (int)out.size());
            return out;
        }, py::arg("values"))
        .def("save_checkpoint", &Engine::save_checkpoint, py::arg("path"))
        .def("load_checkpoint", &Engine::load_checkpoint, py::arg("path"))
        .def("epoch", &Engine::epoch)
        .def("last_kernel_ms", &Engine::last_kernel_ms)
        .def("last_epoch_ms", &Engine::last_epoch_ms)
        .def("last_eval_ms", &Engine::last_eval_ms);

    py::module_ test = m.def_submodule("test", "test");
    py::class_<TestOut, std::shared_ptr<TestOut>>(test, "test_out").def(py::init<>()).def("test", &TestOut::print);
}
