"""Training-object factories: ``load_train_objs()`` of the reference (multigpu.py:122-144, singlegpu.py:132-150) and
the BASELINE.json workloads (toy regressor, wide MLP, GPT-2)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Optional

import numpy as np
import torch
from torch.optim.lr_scheduler import LambdaLR

from . import data as _data
from . import models


@dataclass
class TrainObjs:
    train_set: Any
    model: torch.nn.Module
    optimizer: torch.optim.Optimizer
    test_set: Any
    scheduler: Any

    def __iter__(self):   # unpack like the reference's 5-tuple
        return iter((self.train_set, self.model, self.optimizer, self.test_set, self.scheduler))


def triangular_lr_lambda(steps_per_epoch: int, num_epochs: int = 20):
    """Per-step triangular schedule: 0 -> 1 at 30% of ``num_epochs`` -> 0 (multigpu.py:135-142)."""
    return lambda step: np.interp([step / steps_per_epoch], [0, num_epochs * 0.3, num_epochs], [0, 1, 0])[0]


def vgg_train_objs(world_size: int = 1, synthetic: bool = False, device: Optional[torch.device] = None,
                   data_root: str = "data/cifar10", synthetic_size: int = 50000) -> TrainObjs:
    """The reference's training objects: VGG, SGD(lr .4, momentum .9, wd 5e-4), LambdaLR with the hard-coded
    20-epoch triangle and ``steps_per_epoch`` 98 (single) / 49 (multi, tuned for 2 GPUs - kept, Appendix A)."""
    if synthetic:
        train_set = _data.SyntheticImages(synthetic_size, seed=0) if synthetic_size else None
        test_set = _data.SyntheticImages(min(10000, synthetic_size), seed=1) if synthetic_size else None
    else:
        train_set, test_set = _data.get_training_data(data_root)
    if device is None:
        device = torch.device("cuda") if torch.cuda.is_available() else torch.device("cpu")
    model = models.VGG().to(device)
    optimizer = torch.optim.SGD(model.parameters(), lr=0.4, momentum=0.9, weight_decay=5e-4)
    num_epochs = 20
    steps_per_epoch = 49 if world_size > 1 else 98
    scheduler = LambdaLR(optimizer, triangular_lr_lambda(steps_per_epoch, num_epochs))
    return TrainObjs(train_set, model, optimizer, test_set, scheduler)


def toy_train_objs(size: int = 2048, lr: float = 1e-3, device: Optional[torch.device] = None) -> TrainObjs:
    """Upstream-tutorial workload: Linear(20,1), MSE, SGD(lr 1e-3), 2048 random samples (BASELINE.json config #1)."""
    import sys, os
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from datautils import MyTrainDataset
    train_set = MyTrainDataset(size)
    model = models.ToyLinear()
    if device is not None:
        model = model.to(device)
    optimizer = torch.optim.SGD(model.parameters(), lr=lr)
    return TrainObjs(train_set, model, optimizer, None, None)


def mlp_train_objs(width: int = 8192, depth: int = 8, batch: int = 256, dtype=torch.bfloat16, lr: float = 1e-3,
                   device: Optional[torch.device] = None, dataset_size: int = 0) -> TrainObjs:
    if device is None:
        device = torch.device("cuda") if torch.cuda.is_available() else torch.device("cpu")
    model = models.MLP(width, width, depth, width, dtype=dtype).to(device)
    optimizer = torch.optim.SGD(model.parameters(), lr=lr, momentum=0.9)
    train_set = _data.SyntheticVectors(dataset_size, width, width, dtype=dtype) if dataset_size else None
    return TrainObjs(train_set, model, optimizer, None, None)


def gpt2_train_objs(size: str = "small", dtype=torch.bfloat16, lr: float = 3e-4, device: Optional[torch.device] = None,
                    dataset_size: int = 0, seq_len: int = 1024) -> TrainObjs:
    if device is None:
        device = torch.device("cuda") if torch.cuda.is_available() else torch.device("cpu")
    cfg = getattr(models.GPT2Config, size)()
    model = models.GPT2(cfg, dtype=dtype).to(device)
    optimizer = torch.optim.AdamW(model.parameters(), lr=lr, betas=(0.9, 0.95), weight_decay=0.1)
    train_set = _data.SyntheticTokens(dataset_size, seq_len, cfg.vocab_size) if dataset_size else None
    return TrainObjs(train_set, model, optimizer, None, None)


MODEL_CHOICES = ("vgg", "deepnn", "toy", "mlp", "gpt2", "gpt2-medium", "gpt2-large", "gpt2-xl")


def make(name: str = "vgg", world_size: int = 1, synthetic: bool = False, synthetic_size: int = 50000,
         dtype: Optional[str] = None, device: Optional[torch.device] = None, seq_len: int = 1024):
    """``--model`` of the entrypoints -> (TrainObjs, loss_fn for ``Trainer``).  ``vgg`` is the reference workload
    (multigpu.py:122-144); the others are the BASELINE.json configurations and the reference's dead-code DeepNN
    (singlegpu.py:18-44).  ``dtype``: "fp32" | "bf16" (default: fp32 for the CNNs / toy, bf16 for MLP and GPT-2)."""
    name = name.lower()
    tdt = {None: None, "fp32": torch.float32, "bf16": torch.bfloat16}[dtype]
    if name == "vgg":
        return vgg_train_objs(world_size, synthetic, device, synthetic_size=synthetic_size), None
    if name == "toy":
        return toy_train_objs(device=device), "mse"
    if name == "deepnn":
        if device is None:
            device = torch.device("cuda") if torch.cuda.is_available() else torch.device("cpu")
        if synthetic:
            train_set = _data.SyntheticImages(synthetic_size, seed=0) if synthetic_size else None
            test_set = _data.SyntheticImages(min(10000, synthetic_size), seed=1) if synthetic_size else None
        else:
            train_set, test_set = _data.get_training_data("data/cifar10")
        model = models.DeepNN().to(device)
        optimizer = torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.9)
        return TrainObjs(train_set, model, optimizer, test_set, None), None
    if name == "mlp":
        return mlp_train_objs(dtype=tdt or torch.bfloat16, device=device, dataset_size=synthetic_size), None
    if name.startswith("gpt2"):
        size = {"gpt2": "small", "gpt2-medium": "medium", "gpt2-large": "large", "gpt2-xl": "xl"}[name]
        o = gpt2_train_objs(size, dtype=tdt or torch.bfloat16, device=device, dataset_size=synthetic_size, seq_len=seq_len)
        vocab = o.model.cfg.vocab_size
        from .ops import functional as FX
        return o, (lambda out, tgt: FX.cross_entropy_padded(out, tgt, vocab))
    raise ValueError(f"unknown model {name!r}; choices: {MODEL_CHOICES}")
