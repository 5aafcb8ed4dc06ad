"""A few training steps of the engine without CUDA graphs (so ncu sees plain kernel launches)."""
import os as _os, sys as _sys
_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))

import sys
import torch
from ray_torch_distributed_checkpoint_b200.models import NeuralNetwork
from ray_torch_distributed_checkpoint_b200.parallel.engine import MLPTrainEngine
B = int(sys.argv[1]) if len(sys.argv) > 1 else 18944
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
eng = MLPTrainEngine(NeuralNetwork(), batch_size=B, device="cuda", distributed=False, use_graph=False)
x = torch.randint(0, 256, (B, 784), dtype=torch.uint8, device="cuda"); y = torch.randint(0, 10, (B,), device="cuda")
for _ in range(steps):
    loss = eng.step(x, y)
torch.cuda.synchronize()
print("ok", float(loss), eng.kernel_launches_per_step)
