"""Single-GPU entrypoint - same CLI and call order as the reference's singlegpu.py (:228-263):

    python singlegpu.py TOTAL_EPOCHS SAVE_EVERY [--batch_size 512]
"""
import argparse
import time

import torch

import distributed_data_parallel_experiment_b200 as ddpx
from distributed_data_parallel_experiment_b200 import MiB, Trainer, evaluate, get_model_size
from distributed_data_parallel_experiment_b200.data import prepare_dataloader
from distributed_data_parallel_experiment_b200.models import VGG, DeepNN  # noqa: F401


def load_train_objs(opts=None):
    opts = opts or {}
    name = opts.get("model", "vgg")
    o, loss_fn = ddpx.recipes.make(name, world_size=1, synthetic=bool(opts.get("synthetic", False)),
                                   synthetic_size=int(opts.get("synthetic_size", 50000)), dtype=opts.get("dtype"),
                                   device=(torch.device("cuda") if torch.cuda.is_available() else None) if name == "toy" else None,
                                   seq_len=int(opts.get("seq_len", 1024)))
    opts["_loss_fn"] = loss_fn
    return o.train_set, o.model, o.optimizer, o.test_set, o.scheduler


def main(device, total_epochs, save_every, batch_size, opts=None):
    opts = opts or {}
    dataset, model, optimizer, testdata, scheduler = load_train_objs(opts)
    train_data = prepare_dataloader(dataset, batch_size, distributed=False)
    trainer = Trainer(model, train_data, optimizer, device, save_every, scheduler=scheduler, distributed=False,
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


if __name__ == "__main__":
    parser = argparse.ArgumentParser(description='simple distributed training job')
    parser.add_argument('total_epochs', type=int, help='Total epochs to train the model')
    parser.add_argument('save_every', type=int, help='How often to save a snapshot')
    parser.add_argument('--batch_size', default=512, type=int, help='Input batch size on each device (default: 32)')
    parser.add_argument('--model', default='vgg', choices=list(ddpx.recipes.MODEL_CHOICES), help='extra: model to train')
    parser.add_argument('--dtype', default=None, choices=[None, 'fp32', 'bf16'], help='extra: parameter dtype (MLP / GPT-2)')
    parser.add_argument('--seq-len', dest='seq_len', default=1024, type=int, help='extra: GPT-2 sequence length')
    parser.add_argument('--accum-steps', dest='accum_steps', default=1, type=int, help='extra: micro-batches per optimizer step')
    parser.add_argument('--synthetic', action='store_true', help='extra: CIFAR-shaped random data (no download)')
    parser.add_argument('--synthetic_size', default=50000, type=int)
    args = parser.parse_args()

    device = 0  # shorthand for cuda:0
    main(device, args.total_epochs, args.save_every, args.batch_size,
         {"model": args.model, "synthetic": args.synthetic, "synthetic_size": args.synthetic_size, "dtype": args.dtype,
          "seq_len": args.seq_len, "accum_steps": args.accum_steps})
