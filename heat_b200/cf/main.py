"""Training driver: `python main.py --config <yaml>` (reference: cf_cpu/cf/main.py:19-124).

Same YAML keys and epoch / evaluation schedule.  Differences, all on the GPU-path side:
  * one process per GPU; rank and world size come from the launcher's environment (RANK,
    WORLD_SIZE, LOCAL_RANK -- torchrun) instead of mpi4py; every rank parses the files and takes
    its contiguous user range itself (the reference pickles shards from rank 0, main.py:55-69);
  * evaluation runs fused on the GPU (Engine.evaluate_metrics) and honours the YAML `metrics`
    list; `--dense-eval` takes the reference's route (evaluate0 + numpy metrics) instead;
  * optional `gpu:` section of the YAML (seed, engine, update_mode, max_blocks, warps_per_block, history_mode,
    w0_tail_mode, eval_precision, sync_interval, sync_average, sync_grads, ...); absent = the library defaults
    (engine 0; ranks reconciled by the mean of the replicas every max(65536, 8 x items) samples).
"""
import argparse
import os
import sys
import time

import numpy as np
import torch

if __package__ in (None, ""):
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import cf_c
    import metrics
    import utils
    from behavior_aggregators import AggregatorWeights
    from cf_config import CFConfig
    from datasets import ClickDataset, SubClickDataset
    from models import MatrixFactorization
    from train import Engine
else:
    from . import cf_c, metrics, utils
    from .behavior_aggregators import AggregatorWeights
    from .cf_config import CFConfig
    from .datasets import ClickDataset, SubClickDataset
    from .models import MatrixFactorization
    from .train import Engine


def user_range(num_users, rank, size):
    k, r = divmod(num_users, size)                 # main.py:51-57
    start = rank * k + min(rank, r)
    return start, start + k + (1 if rank < r else 0)


def main(argv=None):
    parser = argparse.ArgumentParser()
    parser.add_argument('--config', type=str, default='./benchmarks/AmazonBooks/MF_CCL/configs/config0.yaml')
    parser.add_argument('--dense-eval', action='store_true', help='evaluate0() + numpy metrics, like the reference')
    parser.add_argument('--checkpoint', type=str, default=None, help='write {tables, grad rows, W0, epoch, LR} here after every '
                        '--checkpoint-interval epochs (rank r writes PATH.rank<r> when WORLD_SIZE > 1)')
    parser.add_argument('--checkpoint-interval', type=int, default=1)
    parser.add_argument('--resume', type=str, default=None, help='continue from a --checkpoint file')
    args = parser.parse_args(argv)
    rank, size = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1))

    config = utils.load_config(args.config)
    dataset_config, model_config = config['dataset_config'], config['model_config']
    gpu = dict(config.get('gpu') or {})
    sync_interval = int(gpu.pop('sync_interval', 0))
    torch.manual_seed(int(model_config.get('seed', 2022)))
    cf_config = CFConfig(emb_dim=model_config['embedding_dim'], num_negs=model_config['num_negs'],
                         max_his=model_config['max_his'], neg_sampler=model_config['neg_sampler'],
                         tile_size=model_config['tile_size'], refresh_interval=model_config['refresh_interval'],
                         l2=model_config['embedding_regularizer'], clip_val=model_config['clip_val'],
                         milestones=model_config['milestones'], l_r=model_config['learning_rate'], **gpu)
    cf_config.seed = int(gpu.get('seed', model_config.get('seed', 2022)))

    test = cf_c.modules.test.test_out()
    print(f'--- Start loading data --- in {rank}')
    sep = dataset_config.get('separator', ' ')
    train_data = ClickDataset(os.path.join(dataset_config['data_dir'], dataset_config['train_data']), separator=sep, config=cf_config)
    test_data = ClickDataset(os.path.join(dataset_config['data_dir'], dataset_config['test_data']), separator=sep, config=cf_config)
    if size > 1:
        start, end = user_range(train_data.num_users, rank, size)
        test.test(f'{start} is the start index'); test.test(f'{end} is the end index')
        train_data = SubClickDataset(train_data, start, end, rank)
        test_data = SubClickDataset(test_data, start, end, rank)
        train_data.update_config(cf_config)
    print(f'--- Finished loading data --- in {rank}')
    cf_config.init_c_instance()

    aggregator_weights = AggregatorWeights(cf_config)
    model = MatrixFactorization(cf_config)
    model.init_c_instance(cf_config)
    engine = Engine(train_data, aggregator_weights, model, cf_config)
    if size > 1:
        import torch.distributed as dist
        dist.init_process_group('gloo')                         # bootstrap only: carries the 128-byte NCCL id
        uid = [cf_c.modules.train.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        engine.c_instance.comm_init(rank, size, uid[0])
        engine.c_instance.set_sync_interval(sync_interval)

    eva_metrics = model_config.get('metrics') or ['Recall(k=20)']
    eval_interval = model_config['eval_interval']
    results = {}
    first_epoch = 0
    suffix = f'.rank{rank}' if size > 1 else ''
    if args.resume:
        engine.load_checkpoint(args.resume + suffix)
        first_epoch = int(engine.c_instance.epoch())
        if rank == 0:
            print(f'--- Resumed from {args.resume} at epoch {first_epoch} ---')
    for epoch in range(first_epoch, model_config['epochs']):
        start_time = time.time()
        epoch_loss = engine.train_one_epoch()
        epoch_time = time.time() - start_time
        if rank == 0:
            n = train_data.click_dataset.shape[0]
            print(f'epoch: {epoch}; loss: {epoch_loss}; epoch_time: {epoch_time}; '
                  f'kernel_ms: {engine.c_instance.last_kernel_ms():.3f}; local samples/s: {n / max(epoch_time, 1e-9):.4g}')
        if args.checkpoint and (epoch + 1) % max(args.checkpoint_interval, 1) == 0:
            engine.save_checkpoint(args.checkpoint + suffix)
        if epoch > 0 and epoch % eval_interval == 0:
            if rank == 0:
                print('--- Start evaluation ---')
            model.eval()
            with torch.no_grad():
                if args.dense_eval:
                    sim_matrix = engine.evaluate0()
                    print(f'sim_matrix shape: {np.shape(sim_matrix)} !!! ')
                    results = metrics.evaluate_metrics(train_data, test_data, sim_matrix, eva_metrics)
                else:
                    results = engine.evaluate_metrics(train_data, test_data, eva_metrics, reduce_over_ranks=size > 1)
                    if rank == 0:
                        print(f'eval_ms: {engine.c_instance.last_eval_ms():.3f}')
    if size > 1:
        import torch.distributed as dist
        dist.destroy_process_group()
    return results


if __name__ == "__main__":
    print('this is main ...')
    main()
