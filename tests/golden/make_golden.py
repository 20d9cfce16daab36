#!/usr/bin/env python
"""Generate golden vectors by EXECUTING THE REFERENCE'S OWN SOURCE FILES.

JAX / distrax / TFP are not installable in this image, so the reference modules
``/root/reference/src/real_nvp.py`` and ``/root/reference/src/BernsteinBijector.py``
are imported unmodified with numpy (float64) standing in for ``jax.numpy``.  Every
line of flow arithmetic and control flow that runs is the reference's; the shim
supplies only the third-party primitives those files call:

  jax.numpy            -> numpy
  jax.nn.leaky_relu    -> where(x>=0, x, 0.01x)         (jax default slope)
  jax.nn.softplus      -> logaddexp(x, 0)
  jax.vmap(f,in_axes=-1)-> map over last axis, stack on axis 0 (jax default out_axes)
  tfd.Beta(a,b).prob   -> scipy.stats.beta.pdf
  tfd.MultivariateNormalDiag(loc, scale_diag).log_prob -> closed form
  distrax.Bijector     -> empty base class

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden.py
Outputs: tests/golden/real_nvp_golden.npz, tests/golden/bernstein_golden.npz
"""
import os
import sys
import types

import numpy as np
import scipy.stats

REF_SRC = "/root/reference/src"
OUT_DIR = os.path.dirname(os.path.abspath(__file__))


def _mod(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def install_shim():
    np.seterr(all="ignore")
    # ---- jax ----
    def leaky_relu(x, negative_slope=0.01):
        return np.where(x >= 0, x, negative_slope * x)

    def softplus(x):
        return np.logaddexp(x, 0.0)

    def sigmoid(x):
        return 1.0 / (1.0 + np.exp(-x))

    def softmax(x, axis=-1):
        e = np.exp(x - np.max(x, axis=axis, keepdims=True))
        return e / e.sum(axis=axis, keepdims=True)

    def vmap(fn, in_axes=0, out_axes=0):
        def mapped(*args):
            n = args[0].shape[in_axes]
            for a in args:
                if a.shape[in_axes] != n:
                    raise ValueError("vmap got inconsistent sizes for array axes to be mapped")
            outs = [fn(*[np.take(a, i, axis=in_axes) for a in args]) for i in range(n)]
            return np.stack(outs, axis=out_axes)

        return mapped

    jnn = _mod("jax.nn", leaky_relu=leaky_relu, softplus=softplus, sigmoid=sigmoid, softmax=softmax)
    jrandom = _mod("jax.random", KeyArray=object, PRNGKey=lambda s: s)
    jax = _mod("jax", nn=jnn, random=jrandom, vmap=vmap, jit=lambda f: f)
    sys.modules["jax.numpy"] = np
    jax.numpy = np

    # ---- tfp ----
    class Beta:
        def __init__(self, concentration1, concentration0):
            self.a = np.asarray(concentration1, dtype=np.float64)
            self.b = np.asarray(concentration0, dtype=np.float64)

        def prob(self, x):
            return scipy.stats.beta.pdf(x, self.a, self.b)

    class MultivariateNormalDiag:
        def __init__(self, loc, scale_diag):
            self.loc, self.scale = np.asarray(loc), np.asarray(scale_diag)

        def log_prob(self, x):
            z = (x - self.loc) / self.scale
            return np.sum(-0.5 * z * z - 0.5 * np.log(2 * np.pi) - np.log(self.scale), axis=-1)

    tfd = types.SimpleNamespace(Beta=Beta, MultivariateNormalDiag=MultivariateNormalDiag)
    tfp_jax = _mod("tensorflow_probability.substrates.jax", distributions=tfd, bijectors=types.SimpleNamespace())
    subs = _mod("tensorflow_probability.substrates", jax=tfp_jax)
    _mod("tensorflow_probability", substrates=subs)

    # ---- distrax ----
    class Bijector:
        def __init__(self, event_ndims_in=0, event_ndims_out=None, **kw):
            self.event_ndims_in = event_ndims_in

    _mod("distrax", Bijector=Bijector)
    _mod("distrax._src")
    _mod("distrax._src.distributions")
    _mod("distrax._src.distributions.distribution", PRNGKey=object)

    # ---- plotting / data deps that the modules import at top level but the hot path never calls ----
    _mod("imageio")
    _mod("seaborn")
    mpl = _mod("matplotlib")
    mpl.pyplot = _mod("matplotlib.pyplot")
    dens = _mod("densities")
    for sub, names in {
        "banana": ["make_dataset_banana"],
        "energies": ["make_dataset_energy_1", "make_dataset_energy_2", "make_dataset_energy_3", "make_dataset_energy_4"],
        "gaussian_blobs": ["make_dataset_gaussian_blobs"],
    }.items():
        setattr(dens, sub, _mod(f"densities.{sub}", **{n: None for n in names}))


def banana(rng, n):
    """src/densities/banana.py:31-41 distribution."""
    x2 = rng.normal(0.0, np.sqrt(3.0), n)
    x1 = rng.normal(x2**2 / 3 - 2, np.sqrt(3.0) / 2)
    return np.stack([x1, x2], axis=-1)


def main():
    install_shim()
    sys.path.insert(0, REF_SRC)
    import BernsteinBijector as ref_bern  # noqa: E402  (the reference file)
    import real_nvp as ref_nvp  # noqa: E402  (the reference file)

    rng = np.random.default_rng(20261104)

    # ------------------------------------------------------------- real_nvp.py
    L = 8
    # init_nvp_chain (real_nvp.py:198-220) needs jax.random; build the same structure by hand
    # with non-zero biases and distinct weights per layer so every term is exercised:
    params, fns_config, flip = [], [], False
    for _ in range(L):
        W1 = rng.normal(size=(1, 4)) * 0.5
        b1 = rng.normal(size=(4,)) * 0.3
        W2 = rng.normal(size=(4, 2)) * 0.5
        b2 = rng.normal(size=(2,)) * 0.2
        params.append([[W1, b1], [W2, b2]])
        fns_config.append((ref_nvp.shift_and_log_scale_fn, flip))
        flip = not flip
    y = np.concatenate([banana(rng, 96), rng.normal(size=(32, 2)) * 3.0])
    z = rng.normal(size=(128, 2))

    logp = ref_nvp.log_prob_nvp_chains(params, fns_config, ref_nvp.log_prob_N01, y)
    samp = ref_nvp.sample_nvp_chain(params, fns_config, lambda N, key: z, 128, None)
    fwd0 = ref_nvp.forward_nvp(params[0], ref_nvp.shift_and_log_scale_fn, z, flip=False)
    fwd1 = ref_nvp.forward_nvp(params[1], ref_nvp.shift_and_log_scale_fn, z, flip=True)
    inv0, ls0 = ref_nvp.inverse_nvp(params[0], ref_nvp.shift_and_log_scale_fn, y, flip=False)
    inv1, ls1 = ref_nvp.inverse_nvp(params[1], ref_nvp.shift_and_log_scale_fn, y, flip=True)
    lp1 = ref_nvp.log_prob_nvp(params[3], ref_nvp.shift_and_log_scale_fn, ref_nvp.log_prob_N01, y, flip=True)
    flat = np.concatenate([t.reshape(-1) for layer in params for (W, b) in layer for t in (W, b)])
    np.savez(
        os.path.join(OUT_DIR, "real_nvp_golden.npz"),
        params_flat=flat, n_layers=L, hidden=4, flips=np.array([f for _, f in fns_config]),
        y=y, z=z, log_prob=logp, sample=samp, fwd0=fwd0, fwd1=fwd1,
        inv0=inv0, ls0=ls0, inv1=inv1, ls1=ls1, log_prob_layer3=lp1,
    )
    print("real_nvp golden: log_prob[:4] =", logp[:4])

    # ------------------------------------------------- BernsteinBijector.py
    out = {}
    theta_demo = np.array([-2.5, 2, 3, -7, -7, -7, -7, -7, 7], dtype=np.float64)  # src/bflow.py:24
    theta_tf = np.array(  # src/bern_flow.py:51-64
        [-1.0744555, -0.6429366, -0.44160688, 0.7950939, 1.9249767, 2.1408765, 2.4256434, 3.1641612, 3.3939004]
    )
    x = np.concatenate([rng.uniform(0, 1, 500), np.array([0.0, 1.0, 1e-9, 1 - 1e-9, 0.5, -0.3, 1.7]),
                        np.linspace(0, 1, 33)])
    out["x"] = x
    for name, th in {"demo": theta_demo, "tf": theta_tf, "m32": rng.normal(size=32), "m64": rng.normal(size=64),
                     "m2": rng.normal(size=2), "m8": rng.normal(size=8)}.items():
        bij = ref_bern.BernsteinBijector(thetas=th)
        yv, ld = bij.forward_and_log_det(x)
        out[f"{name}_theta_raw"] = th
        out[f"{name}_theta_constrained"] = np.asarray(bij.thetas)
        out[f"{name}_y"] = yv
        out[f"{name}_ld"] = ld
    # constrain_thetas on a batch, and the bflow.py:30 double-constrain
    thb = rng.normal(size=(5, 9)) * 2
    out["batch_theta_raw"] = thb
    out["batch_theta_constrained"] = ref_bern.constrain_thetas(thb)
    out["demo_theta_double"] = ref_bern.constrain_thetas(ref_bern.constrain_thetas(theta_demo))
    # per-sample thetas (maf_mwe.py:103-107 style): theta [N, M] against x [N]
    thn = rng.normal(size=(64, 8))
    xn = rng.uniform(0, 1, 64)
    bij = ref_bern.BernsteinBijector(thetas=thn)
    out["ps_theta_raw"], out["ps_x"] = thn, xn
    out["ps_y"], out["ps_ld"] = bij.forward_and_log_det(xn)
    # inverse: thetas [E, M], y [N, E] -> [E, N]
    E = 3
    the = rng.normal(size=(E, 9))
    bij = ref_bern.BernsteinBijector(thetas=the)
    xe = rng.uniform(0.02, 0.98, size=(257, E))
    ye = bij.forward(xe)
    ye[:3] -= 50.0  # below the fitted grid -> clamps to x_fit[0]
    ye[3:6] += 50.0  # above -> clamps to x_fit[-1]
    xinv, zero = bij.inverse_and_log_det(ye)
    out["inv_theta_raw"], out["inv_y"], out["inv_x"], out["inv_logdet"] = the, ye, xinv, np.float64(zero)
    out["inv_x_true"] = xe
    # E = 1 with thetas [M]: y must be [N,1]
    bij = ref_bern.BernsteinBijector(thetas=theta_tf)
    y1 = bij.forward(np.linspace(0, 1, 200))[:, None]
    out["inv1_y"], out["inv1_x"] = y1, bij.inverse(y1)
    np.savez(os.path.join(OUT_DIR, "bernstein_golden.npz"), **out)
    mse = float(np.sum((np.squeeze(out["inv1_x"]) - np.linspace(0, 1, 200)) ** 2))
    print(f"bernstein golden written; reference round-trip SSE (bern_flow.py:79-81 analogue) = {mse:.3e}")


if __name__ == "__main__":
    main()
