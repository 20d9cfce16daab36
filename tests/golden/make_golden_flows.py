#!/usr/bin/env python
"""Golden vectors for the distrax-built flows, produced by EXECUTING THE
REFERENCE'S OWN SCRIPTS:

    /root/reference/src/real_nvp_toy_densities.py   (make_conditioner, make_flow_model, log_prob, sample)
    /root/reference/src/maf_toy_densities.py        (same, RationalQuadraticSpline ±8, 12 bins)
    /root/reference/src/maf_mnist.py                (same, event (28,28,1), 8 bins on [0,1], Uniform base)
    /root/reference/src/bflow.py lines 1-34         (the T1..T4 Bernstein chain, built by the script itself)

are imported / executed unmodified with the numpy stand-ins of ``ref_shim.py`` in place of jax, haiku, distrax
and TFP.  The model wiring — masks, per-layer ``logical_not``, ``3*n+1``, Flatten/Reshape, the ``bijector_fn``
closures, ``Inverse(Chain(layers))``, base distributions, haiku module names — is therefore the reference's
code; the library primitives (RQS, ScalarAffine, Linear, SoftClip ...) are restated in ref_shim.py.

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden_flows.py
Outputs: tests/golden/coupling_toy_golden.npz, coupling_mnist_golden.npz, bflow_golden.npz
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import mnist_params as MP  # noqa: E402
import ref_shim  # noqa: E402

REF_SRC = "/root/reference/src"


def f32(a):
    return np.asarray(a, dtype=np.float32)


def tree64(tree):
    return {m: {k: v.astype(np.float64) for k, v in d.items()} for m, d in tree.items()}


def run_script(mod, tree, x, z_seed, n_sample, event_shape):
    """log_prob.apply / sample.apply exactly as the scripts' loss_fn / main call them."""
    lp = mod.log_prob.apply(tree64(tree), x.astype(np.float64))
    samp = mod.sample.apply(tree64(tree), event_shape, np.random.default_rng(z_seed), n_sample)
    z = ref_shim.LAST_BASE_SAMPLE[0]
    return lp, z, samp


def main():
    ref_shim.install()
    sys.path.insert(0, REF_SRC)
    import maf_mnist as ref_mnist  # noqa: E402  (the reference files)
    import maf_toy_densities as ref_maf  # noqa: E402
    import real_nvp_toy_densities as ref_rnvp  # noqa: E402

    rng = np.random.default_rng(20261105)
    out = {}

    # ---------------------------------------------------------------- real_nvp_toy_densities.py
    x = f32(np.concatenate([rng.normal(size=(200, 2)) * 2.0, rng.normal(size=(56, 2)) * 6.0]))
    init = ref_rnvp.log_prob.init(0, x.astype(np.float64))
    names = sorted(init)
    # the haiku tree the reference's make_flow_model creates: names + shapes are what params_io.py must read
    expect = MP.module_shapes(ref_rnvp.FLOW_NUM_LAYERS, 2, [ref_rnvp.HIDDEN_SIZE] * ref_rnvp.MLP_NUM_LAYERS, 2)
    assert sorted(n for n, _, _ in expect) == names, (names[:5], expect[:5])
    for n, m, k in expect:
        assert init[n]["w"].shape == (m, k) and init[n]["b"].shape == (k,), n
        if not n.startswith("mlp"):
            assert not init[n]["w"].any() and not init[n]["b"].any(), "final Linear must be zero-initialised"
    # K1: identity flow at init -> log_prob = Normal log-density
    lp0 = ref_rnvp.log_prob.apply(init, x.astype(np.float64))
    assert np.allclose(lp0, (-0.5 * x.astype(np.float64) ** 2 - 0.5 * np.log(2 * np.pi)).sum(-1))
    tree = MP.make_tree(101, ref_rnvp.FLOW_NUM_LAYERS, 2, [64, 64], 2, scale=0.3)
    lp, z, samp = run_script(ref_rnvp, tree, x, 7, 128, (2,))
    out.update(rnvp_names=np.array([n for n, _, _ in expect]), rnvp_flat=MP.tree_to_flat(tree, 8, 2, [64, 64], 2),
               rnvp_x=x, rnvp_log_prob=lp, rnvp_z=f32(z), rnvp_sample=None)
    # re-run the sampler on the float32-rounded base noise so that the stored z is exactly what produced the sample
    out["rnvp_sample"] = _sample_with(ref_rnvp, tree, f32(z), (2,))
    print("real_nvp_toy log_prob[:3]", lp[:3])

    # ---------------------------------------------------------------- maf_toy_densities.py
    expect = MP.module_shapes(8, 2, [64, 64], 2 * 37)
    init = ref_maf.log_prob.init(0, x.astype(np.float64))
    assert sorted(init) == sorted(n for n, _, _ in expect)
    assert init["linear_3"]["w"].shape == (64, 74)
    tree = MP.make_tree(202, 8, 2, [64, 64], 74, scale=0.3)
    lp, z, _ = run_script(ref_maf, tree, x, 8, 128, (2,))
    out.update(maf_flat=MP.tree_to_flat(tree, 8, 2, [64, 64], 74), maf_x=x, maf_log_prob=lp, maf_z=f32(z),
               maf_sample=_sample_with(ref_maf, tree, f32(z), (2,)))
    print("maf_toy log_prob[:3]", lp[:3])
    np.savez_compressed(os.path.join(HERE, "coupling_toy_golden.npz"), **out)

    # ---------------------------------------------------------------- maf_mnist.py
    out = {}
    nimg = 24
    img = rng.integers(0, 256, size=(nimg, 28, 28, 1))
    xm = f32((img + rng.uniform(size=img.shape)) / 256.0)  # prepare_data (maf_mnist.py:101-106)
    xm[-1, 3, 4, 0] = 1.25  # outside the Uniform(0,1) support -> log_prob = -inf (linear tail keeps it outside)
    expect = MP.module_shapes(8, 784, [512, 512, 512], 784 * 25)
    tree = MP.make_tree(303, 8, 784, [512, 512, 512], 19600, scale=0.1)
    lp = ref_mnist.log_prob.apply(tree64(tree), xm.astype(np.float64))
    assert lp.shape == (nimg,) and np.isneginf(lp[-1]) and np.isfinite(lp[:-1]).all()
    zu = f32(rng.uniform(size=(4, 28, 28, 1)))
    samp = _sample_with(ref_mnist, tree, zu, (28, 28, 1))
    out.update(mnist_seed=303, mnist_scale=0.1, mnist_checksum=MP.tree_checksum(tree), mnist_x=xm, mnist_log_prob=lp,
               mnist_z=zu, mnist_sample=samp, mnist_names=np.array([n for n, _, _ in expect][:8]))
    print("maf_mnist log_prob[:3]", lp[:3])
    np.savez_compressed(os.path.join(HERE, "coupling_mnist_golden.npz"), **out)

    # ---------------------------------------------------------------- bflow.py:1-34 (+ 82-88, 133-142)
    out = {}
    src = open(os.path.join(REF_SRC, "bflow.py")).read().split("\n")
    prefix = "\n".join(src[:34])  # imports, a/b/theta/alpha/beta, T1..T4, chained_bijectors, bij
    assert "bij = distrax.Inverse(distrax.Chain(chained_bijectors))" in src[33]
    ns = {"__name__": "bflow_prefix"}
    exec(compile(prefix, os.path.join(REF_SRC, "bflow.py"), "exec"), ns)
    import distrax  # the shim

    x1 = np.linspace(-3, 3, 200)
    dist_t = distrax.Transformed(distrax.Normal(loc=0.0, scale=1.0), ns["bij"])  # bflow.py:82-88
    out["bflow_x"] = x1
    out["bflow_log_prob"] = dist_t.log_prob(x1)
    g = np.linspace(-3, 3, 41)
    X = np.stack(np.meshgrid(g, g), axis=-1)  # bflow.py:133-142 on a smaller grid
    blk = distrax.Block(distrax.Inverse(distrax.Chain(ns["chained_bijectors"])), ndims=1)
    out["bflow_X"] = X
    out["bflow_mvn_log_prob"] = distrax.Transformed(distrax.MultivariateNormalDiag(np.zeros(2), np.ones(2)), blk).log_prob(X)
    out["bflow_a"], out["bflow_b"], out["bflow_theta"] = ns["a"], ns["b"], ns["theta"]
    out["bflow_alpha"], out["bflow_beta"] = ns["alpha"], ns["beta"]
    # the individual bijectors of the chain on their own (forward + log-det), for the host-mirror classes
    v = np.linspace(-4, 4, 101)
    for nm, bj in zip(("T4", "T3", "T2", "T1"), ns["chained_bijectors"]):
        dom = np.linspace(0.01, 0.99, 101) if nm == "T3" else v
        yv, ld = bj.forward_and_log_det(dom)
        out[f"bflow_{nm}_x"], out[f"bflow_{nm}_y"], out[f"bflow_{nm}_ld"] = dom, yv, np.broadcast_to(ld, yv.shape)
    total = np.trapz(np.exp(out["bflow_log_prob"]), x1)
    print("bflow log_prob[:3]", out["bflow_log_prob"][:3], " integral over [-3,3] =", total)
    np.savez_compressed(os.path.join(HERE, "bflow_golden.npz"), **out)


def _sample_with(mod, tree, z, event_shape):
    """mod.sample.apply with the base distribution's draw replaced by the given noise: Transformed.sample is
    bijector.forward(base.sample()) — the reference's make_flow_model builds the bijector."""
    import distrax

    orig = distrax.Independent.sample
    distrax.Independent.sample = lambda self, seed, sample_shape=(): z.astype(np.float64)
    try:
        return mod.sample.apply(tree64(tree), event_shape, 0, z.shape[0])
    finally:
        distrax.Independent.sample = orig


if __name__ == "__main__":
    main()
