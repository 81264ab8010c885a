"""The reference network built from the hand-written sm_100a ops (``ops.functional``).

This is the layer-by-layer, autograd-driven form (any topology can be assembled the same way); the
``Engine`` runs the identical kernels from native code with the derivative fusions and a CUDA graph.
"""
import torch
import torch.nn as nn

from ..ops import functional as F200


class B200CNN(nn.Module):
    def __init__(self, pool=False):
        super().__init__()
        self.pool = pool
        self.stride = 1 if pool else 2
        mk = lambda *shape: nn.Parameter(torch.zeros(*shape, dtype=torch.float32))
        self.conv1_w, self.conv1_b = mk(16, 1, 3, 3), mk(16)
        self.conv2_w, self.conv2_b = mk(32, 16, 3, 3), mk(32)
        self.fc1_w, self.fc1_b = mk(200, 1568), mk(200)
        self.fc2_w, self.fc2_b = mk(200, 200), mk(200)
        self.out_w, self.out_b = mk(10, 200), mk(10)

    def param_list(self):
        return [self.conv1_w, self.conv1_b, self.conv2_w, self.conv2_b, self.fc1_w, self.fc1_b, self.fc2_w,
                self.fc2_b, self.out_w, self.out_b]

    def forward(self, x):
        """x: uint8 [N,1,28,28] pixels (normalised in the conv kernel) or float in [0,1]. Returns logits."""
        a = F200.conv2d_act(x, self.conv1_w, self.conv1_b, self.stride, 1, "relu")
        if self.pool:
            a = F200.max_pool2d(a, 2, 2)
        a = F200.conv2d_act(a, self.conv2_w, self.conv2_b, self.stride, 1, "relu")
        if self.pool:
            a = F200.max_pool2d(a, 2, 2)
        h = F200.linear_act(a.flatten(1), self.fc1_w, self.fc1_b, "tanh")
        h = F200.linear_act(h, self.fc2_w, self.fc2_b, "tanh")
        return F200.linear_act(h, self.out_w, self.out_b, "none")

    def loss_sum(self, x, labels):
        loss, _probs, _metrics = F200.softmax_xent(self.forward(x), labels)
        return loss
