"""Multi-GPU entrypoint - same CLI and call order as the reference's multigpu.py (:224-263):

    python multigpu.py TOTAL_EPOCHS SAVE_EVERY [--batch_size 512]

One process per local GPU (``spawn``), ``ddp_setup`` -> ``load_train_objs`` -> ``prepare_dataloader`` ->
``Trainer(...).train()`` -> timing / model-size / accuracy prints -> ``destroy_process_group``.  Extra, optional flags
select synthetic data, other models and the communication backend; defaults reproduce the reference.
"""
import argparse
import os
import time

import torch

import distributed_data_parallel_experiment_b200 as ddpx
from distributed_data_parallel_experiment_b200 import MiB, Trainer, evaluate, get_model_size
from distributed_data_parallel_experiment_b200.data import prepare_dataloader
from distributed_data_parallel_experiment_b200.launch import ddp_setup, destroy_process_group
from distributed_data_parallel_experiment_b200.models import VGG  # noqa: F401  (re-exported like the reference)


def _apply_runtime_opts(opts):
    """Optional knobs -> runtime configuration (before the data-parallel wrapper is built)."""
    cfg = ddpx.config.get()
    if opts.get("bucket_cap_mb") is not None:
        cfg.bucket_cap_mb = float(opts["bucket_cap_mb"])
    if opts.get("allreduce_algo"):
        cfg.allreduce_algo = opts["allreduce_algo"]


def load_train_objs(opts=None):
    world = torch.distributed.get_world_size() if torch.distributed.is_initialized() else 1
    opts = opts or {}
    model_name = opts.get("model", "vgg")
    device = torch.device("cuda") if ddpx.launch.backend() != "gloo" and torch.cuda.is_available() else torch.device("cpu")
    o, loss_fn = ddpx.recipes.make(model_name, world_size=max(world, 2), synthetic=bool(opts.get("synthetic", False)),
                                   synthetic_size=int(opts.get("synthetic_size", 50000)), dtype=opts.get("dtype"),
                                   device=None if model_name == "toy" else device, seq_len=int(opts.get("seq_len", 1024)))
    opts["_loss_fn"] = loss_fn
    if opts.get("sync_bn"):
        ddpx.nn.convert_sync_batchnorm(o.model)      # the option the reference keeps commented out (multigpu.py:127)
    return o.train_set, o.model, o.optimizer, o.test_set, o.scheduler


def main(rank: int, world_size: int, save_every: int, total_epochs: int, batch_size: int, opts=None):
    opts = opts or {}
    ddp_setup(rank, world_size, backend=opts.get("backend"))
    _apply_runtime_opts(opts)
    dataset, model, optimizer, testdata, scheduler = load_train_objs(opts)
    train_data = prepare_dataloader(dataset, batch_size)
    trainer = Trainer(model, train_data, optimizer, rank, save_every, scheduler=scheduler,
                      loss_fn=opts.get("_loss_fn"), accum_steps=int(opts.get("accum_steps", 1)))

    start_time = time.time()
    trainer.train(total_epochs)
    end_time = time.time()

    training_time = end_time - start_time
    print(f"Total training time: {training_time:.2f} seconds")

    fp32_model_size = get_model_size(model)
    print(f"fp32 model has size={fp32_model_size/MiB:.2f} MiB")

    if testdata is not None:
        from torch.utils.data import DataLoader
        test_data = DataLoader(testdata, batch_size=512, shuffle=False, pin_memory=torch.cuda.is_available(), num_workers=0)
        fp32_model_accuracy = evaluate(model, test_data)
        print(f"fp32 model has accuracy={fp32_model_accuracy:.2f}%")

    destroy_process_group()


if __name__ == "__main__":
    parser = argparse.ArgumentParser(description='simple distributed training job')
    parser.add_argument('total_epochs', type=int, help='Total epochs to train the model')
    parser.add_argument('save_every', type=int, help='How often to save a snapshot')
    parser.add_argument('--batch_size', default=512, type=int, help='Input batch size on each device (default: 32)')
    parser.add_argument('--model', default='vgg', choices=list(ddpx.recipes.MODEL_CHOICES), help='extra: model to train')
    parser.add_argument('--dtype', default=None, choices=[None, 'fp32', 'bf16'], help='extra: parameter dtype (MLP / GPT-2)')
    parser.add_argument('--seq-len', dest='seq_len', default=1024, type=int, help='extra: GPT-2 sequence length')
    parser.add_argument('--bucket-cap-mb', dest='bucket_cap_mb', default=None, type=float, help='extra: gradient bucket cap')
    parser.add_argument('--allreduce-algo', dest='allreduce_algo', default=None,
                        choices=[None, 'auto', 'oneshot', 'twoshot', 'nvls', 'nccl'], help='extra: all-reduce algorithm')
    parser.add_argument('--sync-bn', dest='sync_bn', action='store_true', help='extra: SyncBatchNorm (statistics over all ranks)')
    parser.add_argument('--accum-steps', dest='accum_steps', default=1, type=int, help='extra: micro-batches per optimizer step (no_sync)')
    parser.add_argument('--synthetic', action='store_true', help='extra: CIFAR-shaped random data (no download)')
    parser.add_argument('--synthetic_size', default=50000, type=int)
    parser.add_argument('--backend', default=None, choices=[None, 'auto', 'p2p', 'nccl', 'gloo'], help='extra: gradient backend')
    parser.add_argument('--nprocs', default=None, type=int, help='extra: number of processes (default: all local GPUs)')
    args = parser.parse_args()

    world_size = args.nprocs or (torch.cuda.device_count() if torch.cuda.is_available() else 2)
    opts = {"model": args.model, "synthetic": args.synthetic, "synthetic_size": args.synthetic_size, "backend": args.backend,
            "dtype": args.dtype, "seq_len": args.seq_len, "bucket_cap_mb": args.bucket_cap_mb,
            "allreduce_algo": args.allreduce_algo, "accum_steps": args.accum_steps, "sync_bn": args.sync_bn}
    ddpx.spawn(main, args=(world_size, args.save_every, args.total_epochs, args.batch_size, opts), nprocs=world_size)
