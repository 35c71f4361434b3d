import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
import test_fused_step_gpu as T
import tfde_b200 as tf
from tfde_b200.parallel.strategy import Strategy
x, y = T._data(512)
for kind in ["adam", "sgd"]:
    ms = {}
    for dev in ["cpu", "cuda"]:
        with Strategy(device=dev).scope():
            ms[dev] = tf.models.build_and_compile_cnn_model(optimizer=T._opt(kind), seed=7)
    for i in range(8):
        for dev in ["cpu", "cuda"]:
            ms[dev].train_on_batch(x[i*64:(i+1)*64], y[i*64:(i+1)*64])
        lc, lg = ms["cpu"].last_step_loss(), ms["cuda"].last_step_loss()
        errs = []
        for a, b, v in zip(ms["cuda"].get_weights(), ms["cpu"].get_weights(), ms["cuda"].variables):
            errs.append((v.name, float(np.abs(a - b).max()), float(np.abs(b).max())))
        print(kind, i, lc, lg, " ".join(f"{n}:{e:.2e}/{m:.2e}" for n, e, m in errs), flush=True)
    from tfde_b200.ops import native
    print(native.hparams_get(ms["cuda"].optimizer.hparams))
