"""Top-1 accuracy evaluation, same contract as the reference's ``evaluate`` (multigpu.py:180-205)."""
from __future__ import annotations

import torch
from torch import nn


@torch.inference_mode()
def evaluate(model: nn.Module, dataflow, device=None, progress: bool = True) -> float:
    """Runs ``model`` in eval mode over ``dataflow`` and returns accuracy in percent.

    Like the reference, the model is left in eval mode and the correct-count is accumulated on the device with a
    single host read at the end (multigpu.py:203-205).
    """
    model.eval()
    if device is None:
        try:
            device = next(model.parameters()).device
        except StopIteration:
            device = torch.device("cpu")
    it = dataflow
    if progress:
        try:
            from tqdm.auto import tqdm
            it = tqdm(dataflow, desc="eval", leave=False)
        except Exception:
            it = dataflow
    num_samples = 0
    num_correct = 0
    for inputs, targets in it:
        inputs = inputs.to(device, non_blocking=True)
        targets = targets.to(device, non_blocking=True)
        outputs = model(inputs)
        outputs = outputs.argmax(dim=1)
        num_samples += targets.size(0)
        num_correct += (outputs == targets).sum()
    return (num_correct / num_samples * 100).item()
