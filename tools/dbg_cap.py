import os, sys, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import tfde_b200 as tf
from tfde_b200.ops import native, mixed
from tfde_b200.parallel.strategy import Strategy
L = native.load()
L.tfde_capture_status.argtypes=[ctypes.c_void_p]; L.tfde_capture_status.restype=ctypes.c_int
orig_check = native.check
def check(code, what, kernels=1):
    st = L.tfde_capture_status(native.stream_ptr())
    print(f"   [{what}] rc={code} capture_status={st}", flush=True)
    return orig_check(code, what, kernels)
native.check = check
import tfde_b200.ops.mixed as mx
# replicate test_mixed's training test first
tf.keras.mixed_precision.set_global_policy("mixed_bfloat16")
g = torch.Generator().manual_seed(0)
x = torch.rand(256, 28, 28, 1, generator=g); y = torch.randint(0, 10, (256,), generator=g)
with Strategy(device="cuda").scope():
    m = tf.models.build_and_compile_cnn_model(learning_rate=0.05, seed=2)
for i in range(5):
    print("mixed step", i, flush=True)
    m.train_on_batch(x, y); m.last_step_loss()
tf.keras.mixed_precision.set_global_policy("float32")
print("=== smoke", flush=True)
import __graft_entry__ as ge
ge.smoke()
