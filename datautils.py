"""Toy dataset of the upstream tutorial the reference was forked from (`# from datautils import MyTrainDataset`,
multigpu.py:4): ``size`` samples of ``(rand(20), rand(1))``.  BASELINE.json config #1 trains Linear(20,1)+MSE on it."""
import torch
from torch.utils.data import Dataset


class MyTrainDataset(Dataset):
    def __init__(self, size):
        self.size = size
        self.data = [(torch.rand(20), torch.rand(1)) for _ in range(size)]

    def __len__(self):
        return self.size

    def __getitem__(self, index):
        return self.data[index]
