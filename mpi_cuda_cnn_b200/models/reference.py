"""PyTorch reference model: oracle for the kernels and the NCCL-baseline network.

Topology, layouts and activations follow /root/reference/cnn.c:416-428 (conv 3x3 s2 p1 + ReLU twice,
two 200-unit tanh layers, 10-way softmax); conv weights are OIHW and linear weights [out][in], which
is how the reference allocates them (cnn.c:333, :116-121), so parameters move between the CPU
engine, the GPU engine and this model without permutation.
"""
import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F


class RefCNN(nn.Module):
    def __init__(self, pool=False, dtype=torch.float64):
        super().__init__()
        self.pool = pool
        s = 1 if pool else 2
        self.conv1 = nn.Conv2d(1, 16, 3, stride=s, padding=1, dtype=dtype)
        self.conv2 = nn.Conv2d(16, 32, 3, stride=s, padding=1, dtype=dtype)
        self.fc1 = nn.Linear(32 * 7 * 7, 200, dtype=dtype)
        self.fc2 = nn.Linear(200, 200, dtype=dtype)
        self.out = nn.Linear(200, 10, dtype=dtype)

    def param_list(self):
        """Layer order, weights then biases per layer — the dense order of both native engines."""
        return [self.conv1.weight, self.conv1.bias, self.conv2.weight, self.conv2.bias, self.fc1.weight,
                self.fc1.bias, self.fc2.weight, self.fc2.bias, self.out.weight, self.out.bias]

    def features(self, x):
        """All activations, for layer-by-layer comparison.  x: [N, 1, 28, 28] in [0, 1]."""
        a1 = F.relu(self.conv1(x))
        if self.pool:
            a1 = F.max_pool2d(a1, 2, 2)
        a2 = F.relu(self.conv2(a1))
        if self.pool:
            a2 = F.max_pool2d(a2, 2, 2)
        h1 = torch.tanh(self.fc1(a2.flatten(1)))
        h2 = torch.tanh(self.fc2(h1))
        logits = self.out(h2)
        return a1, a2, h1, h2, logits

    def forward(self, x):
        return self.features(x)[-1]

    def loss_sum(self, x, labels):
        """Sum (not mean) of cross-entropy over the batch: its gradient w.r.t. the logits is
        softmax - onehot per sample, the reference's errors (cnn.c:285-286), summed like u_weights."""
        return F.cross_entropy(self.forward(x), labels.long(), reduction="sum")

    @staticmethod
    def mse_metric(logits, labels):
        """The reference's printed 'error': mean over samples of mean_j (p_j - y_j)^2 (cnn.c:275-282)."""
        p = torch.softmax(logits, dim=1)
        y = F.one_hot(labels.long(), p.shape[1]).to(p.dtype)
        return ((p - y) ** 2).mean(dim=1).mean()


def load_dense_params(model, dense):
    dense = torch.as_tensor(np.asarray(dense))
    o = 0
    with torch.no_grad():
        for p in model.param_list():
            n = p.numel()
            p.copy_(dense[o:o + n].reshape(p.shape).to(p.dtype))
            o += n
    assert o == dense.numel()
    return model


def dense_params(model):
    return torch.cat([p.detach().reshape(-1) for p in model.param_list()]).cpu().numpy()


def dense_grads(model):
    return torch.cat([p.grad.detach().reshape(-1) for p in model.param_list()]).cpu().numpy()
