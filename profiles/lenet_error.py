"""Error of the LeNet logits against a float64 numpy evaluation of the same network, for the
tensor-core kernels, the fp32 CUDA-core kernels and the CPU oracle (fp32, sequential sums).
Run on the GPU box:  python profiles/lenet_error.py [n_images]
Each kernel variant runs in its own process (AG2_LENET_SIMT is read once per process)."""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def images(n):
    rng = np.random.default_rng(17)
    base = rng.integers(0, 256, size=(n, 60, 60, 3), dtype=np.uint8)
    base[rng.random(base.shape) < 0.7] = 0
    return base


def lenet_f64(w, imgs):
    c1w, c1b, c2w, c2b, i1w, i1b, i2w, i2b = [np.asarray(w[k], np.float64) for k in ("conv1_w", "conv1_b", "conv2_w", "conv2_b", "ip1_w", "ip1_b", "ip2_w", "ip2_b")]
    x = imgs.astype(np.float64).transpose(0, 3, 1, 2)  # n,3,60,60

    def conv(x, wt, b, co, ci):
        wt = wt.reshape(co, ci, 5, 5)
        win = np.lib.stride_tricks.sliding_window_view(x, (5, 5), axis=(2, 3))  # n,ci,H,W,5,5
        return np.einsum("nchwyx,ocyx->nohw", win, wt, optimize=True) + b[None, :, None, None]

    def pool(x):
        n, c, h, w_ = x.shape
        return x.reshape(n, c, h // 2, 2, w_ // 2, 2).max(axis=(3, 5))

    x = pool(conv(x, c1w, c1b, 20, 3))
    x = pool(conv(x, c2w, c2b, 50, 20))
    h = np.maximum(x.reshape(len(x), -1) @ i1w.reshape(500, 7200).T + i1b, 0)
    return h @ i2w.reshape(2, 500).T + i2b


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 43
    if len(sys.argv) > 2:  # child: print the GPU logits
        from agile_grasp2_b200 import capi
        from common import make_params
        from oracle import binding as ob
        w = ob.xavier_weights(11, bias_scale=0.2)
        g = capi.Context(make_params(capi.MODE_PREDICTION))
        g.set_lenet_weights(w)
        rep = int(sys.argv[2])
        idx = np.arange(rep) % n
        gl = g.classify_images(np.ascontiguousarray(images(n)[idx]))[:n]
        np.save(sys.argv[3], gl)
        return
    from common import make_params
    from agile_grasp2_b200 import capi
    from oracle import binding as ob
    w = ob.xavier_weights(11, bias_scale=0.2)
    imgs = images(n)
    ref = lenet_f64(w, imgs)
    o = ob.context(make_params(capi.MODE_PREDICTION))
    o.set_lenet_weights(w)
    ol = o.classify_images(imgs)
    rel = lambda a: (np.abs(a - ref) / np.abs(ref)).max()
    sgn = lambda a: float(np.mean((np.abs(a) - np.abs(ref)) / np.abs(ref)))
    print("logit magnitude: median %.1f max %.1f" % (np.median(np.abs(ref)), np.abs(ref).max()))
    print("%-34s max rel %.2e   mean signed rel (|x| vs |ref|) %+.2e" % ("oracle (fp32 CPU)", rel(ol), sgn(ol)))
    for mode, name in (("1", "CUDA cores fp32"), ("2", "tensor conv2 + CUDA-core ip1"), ("0", "tensor conv1 + conv2 + ip1")):
        for rep in (n, 4500):
            out = "/tmp/lenet_err_%s_%d.npy" % (mode, rep)
            env = dict(os.environ, AG2_LENET_SIMT=mode)
            subprocess.run([sys.executable, os.path.abspath(__file__), str(n), str(rep), out], env=env, check=True)
            gl = np.load(out)
            print("%-34s max rel %.2e   mean signed rel (|x| vs |ref|) %+.2e" % ("%s, batch %d" % (name, rep), rel(gl), sgn(gl)))


if __name__ == "__main__":
    main()
