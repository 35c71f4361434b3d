"""Host-side cost breakdown of Sequential.train_on_batch in steady state (graph replay)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import tfde_b200 as tf

m = tf.models.build_and_compile_cnn_model(seed=1)
g = torch.Generator().manual_seed(0)
xs = [torch.randint(0, 256, (64, 28, 28, 1), generator=g, dtype=torch.uint8).pin_memory() for _ in range(8)]
ys = [torch.randint(0, 10, (64,), generator=g).pin_memory() for _ in range(8)]
for i in range(20):
    m.train_on_batch(xs[i % 8], ys[i % 8])
m.sync()
N = 2000
# (a) launch only, no reads, no sync
t0 = time.perf_counter()
for i in range(N):
    m.train_on_batch(xs[i % 8], ys[i % 8])
t1 = time.perf_counter()
m.sync()
t2 = time.perf_counter()
print(f"train_on_batch host time {1e6*(t1-t0)/N:.1f} us/step; incl. drain {1e6*(t2-t0)/N:.1f} us/step")
# (b) with lag-1 loss read
t0 = time.perf_counter()
for i in range(N):
    m.train_on_batch(xs[i % 8], ys[i % 8])
    if i:
        m.last_step_loss(lag=1)
m.last_step_loss()
t1 = time.perf_counter()
print(f"with lag-1 loss read {1e6*(t1-t0)/N:.1f} us/step")
# (c) pieces
st = [v for v in m._graphs.values() if hasattr(v, "slots")][0]
slot = st.slots[0]
t0 = time.perf_counter()
for i in range(N):
    slot.host_x.copy_(xs[i % 8]); slot.host_y.copy_(ys[i % 8])
t1 = time.perf_counter()
print(f"two host staging copies {1e6*(t1-t0)/N:.1f} us")
t0 = time.perf_counter()
for i in range(N):
    with torch.cuda.stream(m._stream):
        pass
t1 = time.perf_counter()
print(f"stream context {1e6*(t1-t0)/N:.1f} us")
torch.cuda.synchronize()
with torch.cuda.stream(m._stream):
    t0 = time.perf_counter()
    for i in range(N):
        slot.graph.replay()
    t1 = time.perf_counter()
    m._stream.synchronize()
    t2 = time.perf_counter()
print(f"graph.replay host {1e6*(t1-t0)/N:.1f} us; GPU-bound rate {1e6*(t2-t0)/N:.1f} us/step (H2D+step+D2H, warm L2)")
ev = torch.cuda.Event()
t0 = time.perf_counter()
for i in range(N):
    ev.record(m._stream)
t1 = time.perf_counter()
print(f"event.record {1e6*(t1-t0)/N:.1f} us")
g2 = m.device_step_graph(st, 64)
with torch.cuda.stream(m._stream):
    m._stream.synchronize()
    t0 = time.perf_counter()
    for i in range(N):
        g2.replay()
    m._stream.synchronize()
    t2 = time.perf_counter()
print(f"device-only graph GPU-bound rate {1e6*(t2-t0)/N:.1f} us/step (warm L2)")
