import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench
from deepsdf_b200.trainer import SceneTrainer
dev = torch.device("cuda:0")
gen = torch.Generator(device=dev).manual_seed(2)
ids = torch.arange(8) * 3
batch = bench.make_batch(gen, dev, ids, bench.SCENES_TOTAL)
out = []
for key_seed in (0, 0, 5):
    tr = SceneTrainer(bench.SPECS, bench.SCENES_TOTAL, dev, seed=0)
    tr.seed = key_seed            # only the dropout key depends on it after construction
    for _ in range(5):
        loss = tr.step(batch, ids.to(dev), epoch=1)
    out.append(float(loss[0]))
print("same key twice:", out[0], out[1], " other key:", out[2])
