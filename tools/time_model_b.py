"""Step time of Model B (mnist_keras_distributed.py:79-109) through train_on_batch: host inputs (graph with staging
slots) and device inputs (step graph), plus the eager per-kernel launch count."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import tfde_b200 as tf
from tfde_b200.ops import native

B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
m = tf.models.build_and_compile_model_b(learning_rate=0.01, seed=1)
g = torch.Generator().manual_seed(0)
x = torch.rand(B, 784, generator=g).pin_memory()
y = torch.randint(0, 10, (B,), generator=g).pin_memory()
c0 = native.launch_count()
m.train_on_batch(x, y); m.sync()
print("kernel entry points per eager step:", native.launch_count() - c0)
for mode, (xx, yy) in {"host": (x, y), "device": (x.cuda(), y.cuda())}.items():
    for _ in range(10):
        m.train_on_batch(xx, yy)
    m.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); m.record_event(e0)
    N = 200
    for _ in range(N):
        m.train_on_batch(xx, yy)
    m.record_event(e1); m.sync()
    print(f"{mode}: wall {1e6 * (time.perf_counter() - t0) / N:.1f} us/step, device {1e3 * e0.elapsed_time(e1) / N:.1f} us/step, graph={m.graph_captured}")
