"""The model and data both variants score (reference examples/features/torch_batch_process_core_api_comparison/model.py)."""
import torch
from torch import nn

N_RECORDS = 1000


def get_model() -> nn.Module:
    torch.manual_seed(11)
    return nn.Sequential(nn.Flatten(), nn.Linear(784, 32), nn.ReLU(), nn.Linear(32, 10)).eval()


class Images(torch.utils.data.Dataset):
    def __init__(self) -> None:
        self.x = torch.randn(N_RECORDS, 1, 28, 28, generator=torch.Generator().manual_seed(3))

    def __len__(self) -> int:
        return len(self.x)

    def __getitem__(self, i: int) -> torch.Tensor:
        return self.x[i]
