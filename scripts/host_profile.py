"""cProfile of the host side of one training step (GPU box)."""
import cProfile, pstats, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import bench
from gnnepcsaft_b200 import create_model, graph, ops, synth
from gnnepcsaft_b200.train import Trainer
dev = torch.device("cuda", 0)
ops.set_matmul_precision("tf32")
cfg = bench.WORKLOADS["cfg2"]
graphs = bench.make_dataset(0, 4096)
store = graph.DeviceGraphStore(synth.pack(graphs), dev)
model = create_model(dict(cfg), store.calc_deg()).to(dev).train()
tr = Trainer(model)
ids = np.arange(4096)
for _ in range(3):
    tr.step(store.collate(ids))
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(5):
    tr.step(store.collate(ids))
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(28)
