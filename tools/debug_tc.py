import os, sys
os.environ["TFDE_FUSED_TC"] = "1"; os.environ["TFDE_FUSED_TC_FORCE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
import test_fused_step_gpu as T
import tfde_b200 as tf
from tfde_b200.parallel.strategy import Strategy
x, y = T._data(512, seed=21)
ms = {}
for dev in ["cpu", "cuda"]:
    with Strategy(device=dev).scope():
        ms[dev] = tf.models.build_and_compile_cnn_model(optimizer=T._opt("sgd"), seed=7)
for i in range(3):
    for dev in ["cpu", "cuda"]:
        ms[dev].train_on_batch(x[i*64:(i+1)*64], y[i*64:(i+1)*64])
    lc, lg = ms["cpu"].last_step_loss(), ms["cuda"].last_step_loss()
    gc, gg = ms["cpu"].arena.g.numpy(), ms["cuda"].arena.g.cpu().numpy()
    errs = []
    for v in ms["cuda"].variables:
        a, b = gg[v.offset:v.offset+v.size], gc[v.offset:v.offset+v.size]
        errs.append((v.name, float(np.abs(a-b).max()), float(np.abs(b).max())))
    print(i, lc, lg, " ".join(f"{n}:{e:.2e}/{m:.2e}" for n, e, m in errs), flush=True)
v = [v for v in ms["cuda"].variables if v.name == "dense/kernel"][0]
a = gg[v.offset:v.offset+v.size].reshape(169, 32, 64); b = gc[v.offset:v.offset+v.size].reshape(169, 32, 64)
d = np.abs(a - b)
print("dense/kernel grad err by pixel (max):", d.max(axis=(1, 2))[:8], "by channel:", d.max(axis=(0, 2))[:8], "by unit:", d.max(axis=(0, 1))[:8])
