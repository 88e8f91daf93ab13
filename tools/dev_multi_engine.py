"""12 objects (two engines of one DeAOTInferEngine) at 480p: frames/s with the groups' read paths as graphs."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench as B
from rmem_ocu_b200 import ops
from rmem_ocu_b200.engine import DeAOTInferEngine
dev = torch.device("cuda", 0)
model = B.build_model()
imgs, _ = B.make_clip(60, 1000)
lab = torch.zeros(1, 1, B.H, B.W, dtype=torch.int32)
for o in range(1, 13):
    lab[:, :, 20 + 30 * (o - 1):20 + 30 * o, 40 + 60 * (o - 1):120 + 60 * (o - 1)] = o
eng = DeAOTInferEngine(model, 0, long_term_mem_gap=5)
d = [im.to(dev) for im in imgs]
with torch.no_grad():
    eng.add_reference_frame(d[0], lab.to(dev), obj_nums=[12], frame_step=0)
    def step(f):
        lg = eng.match_propogate_one_frame(d[f], output_size=(B.H, B.W))
        eng.update_memory(ops.argmax_label(lg))
    for f in range(1, 30): step(f)
    torch.cuda.synchronize(); t = time.perf_counter()
    for f in range(30, 60): step(f)
    torch.cuda.synchronize(); dt = time.perf_counter() - t
e = eng.aot_engines
print("engines %d, %.0f frames/s (12 objects), read graphs per engine: %s" % (len(e), 30 / dt, [sum(1 for k in x._graphs if k[0] == "read") for x in e]))
