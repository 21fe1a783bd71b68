"""The 784 -> hidden -> 10 MLP of the reference's dist-mnist workload
(examples/v1/dist-mnist/dist_mnist.py:166-194) and the small CNN of its Keras NCCL example
(examples/v1/distribution_strategy/keras-API/multi_worker_strategy-with-keras.py:42-60)."""
import torch.nn as nn


class MnistMLP(nn.Module):
    def __init__(self, hidden: int = 100):
        super().__init__()
        self.hid = nn.Linear(784, hidden)
        self.sm = nn.Linear(hidden, 10)

    def forward(self, x):
        return self.sm(self.hid(x.flatten(1)).relu())


class MnistCNN(nn.Module):
    """conv32-pool-conv64-pool-conv64-flatten-dense64-dense10 (93,322 parameters)."""

    def __init__(self):
        super().__init__()
        self.net = nn.Sequential(
            nn.Conv2d(1, 32, 3), nn.ReLU(), nn.MaxPool2d(2), nn.Conv2d(32, 64, 3), nn.ReLU(),
            nn.MaxPool2d(2), nn.Conv2d(64, 64, 3), nn.ReLU(), nn.Flatten(),
            nn.Linear(64 * 3 * 3, 64), nn.ReLU(), nn.Linear(64, 10))

    def forward(self, x):
        return self.net(x)
