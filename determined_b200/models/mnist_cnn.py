"""The MNIST tutorial CNN — BASELINE config #1 (reference ``examples/tutorials/mnist_pytorch/model.py:17-35``
architecture: conv(1→n1,3) relu conv(n1→n2,3) relu maxpool2 dropout flatten fc(…→128) relu dropout fc(128→10)
log_softmax; 1 199 882 parameters at n1=32, n2=64)."""
import torch
import torch.nn as nn
import torch.nn.functional as F


class MNISTNet(nn.Module):
    def __init__(self, n_filters1: int = 32, n_filters2: int = 64, dropout1: float = 0.25, dropout2: float = 0.5) -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(1, n_filters1, 3, 1)
        self.conv2 = nn.Conv2d(n_filters1, n_filters2, 3, 1)
        self.drop1 = nn.Dropout(dropout1)
        self.drop2 = nn.Dropout(dropout2)
        self.fc1 = nn.Linear(144 * n_filters2, 128)
        self.fc2 = nn.Linear(128, 10)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = F.relu(self.conv1(x))
        x = F.max_pool2d(F.relu(self.conv2(x)), 2)
        x = self.drop1(x).flatten(1)
        x = self.drop2(F.relu(self.fc1(x)))
        return F.log_softmax(self.fc2(x), dim=1)
