"""Host-side parameter interchange (normalizing_flows_b200/params_io.py): the
haiku-named tree and the nested-list format map onto the flat buffer the kernels read, in the
oracle's own flatten order."""
import numpy as np
import pytest
import torch

from normalizing_flows_b200 import params_io as io
from oracle import flows_oracle as O


def _tree_from_layers(params):
    names = io.haiku_module_names(len(params), len(params[0]) - 1)
    return {mod: {"w": W.numpy(), "b": b.numpy()} for layer, mods in zip(params, names) for (W, b), mod in zip(layer, mods)}


def test_haiku_names_follow_creation_order():
    names = io.haiku_module_names(3, 2)
    assert names[0] == ["mlp/~/linear_0", "mlp/~/linear_1", "linear"]
    assert names[1] == ["mlp_1/~/linear_0", "mlp_1/~/linear_1", "linear_1"]
    assert names[2][-1] == "linear_2"


@pytest.mark.parametrize("spec", [O.RQS_TOY, O.REAL_NVP_TOY, O.CouplingSpec(16, 3, [8, 8, 8], "rqs", 8, 0.0, 1.0, "uniform")])
def test_haiku_tree_and_npz_round_trip(tmp_path, spec):
    params = O.init_coupling_params(spec, seed=2, dtype=torch.float32)
    flat_ref = O.flatten_layers(params)
    tree = _tree_from_layers(params)
    args = (spec.num_layers, spec.event_dim, spec.hidden, spec.cond_out)
    flat = io.haiku_to_flat(tree, *args)
    assert flat.dtype == torch.float32 and torch.equal(flat, flat_ref)
    back = io.flat_to_haiku(flat, *args)
    assert sorted(back) == sorted(tree)
    for mod in tree:
        assert np.array_equal(back[mod]["w"], tree[mod]["w"]) and np.array_equal(back[mod]["b"], tree[mod]["b"])
    path = str(tmp_path / "params.npz")
    io.save_haiku_npz(path, flat, *args)
    with np.load(path) as z:
        assert "mlp/~/linear_0/w" in z.files and f"linear_{spec.num_layers - 1}/b" in z.files
    assert torch.equal(io.load_haiku_npz(path, *args), flat_ref)


def test_bad_trees_fail_loudly(tmp_path):
    spec = O.REAL_NVP_TOY
    params = O.init_coupling_params(spec, seed=2, dtype=torch.float32)
    args = (spec.num_layers, spec.event_dim, spec.hidden, spec.cond_out)
    tree = _tree_from_layers(params)
    missing = dict(tree)
    del missing["mlp_3/~/linear_1"]
    with pytest.raises(KeyError):
        io.haiku_to_flat(missing, *args)
    wrong = dict(tree)
    wrong["linear"] = {"w": np.zeros((64, 3), np.float32), "b": np.zeros(3, np.float32)}
    with pytest.raises(ValueError):
        io.haiku_to_flat(wrong, *args)
    with pytest.raises(ValueError):
        io.flat_to_haiku(torch.zeros(5), *args)
    path = str(tmp_path / "bad.npz")
    np.savez(path, weights=np.zeros(3))
    with pytest.raises(ValueError):
        io.load_haiku_npz(path, *args)


def test_nested_lists_of_the_hand_rolled_real_nvp():
    """params[layer][sublayer] = [W, b] (src/real_nvp.py:188-195): 8 layers of [1,4,2] = 144 floats."""
    params, _flips = O.init_nvp_chain_like(8, hidden=4, seed=534, dtype=torch.float32)
    nested = [[[W.numpy(), b.numpy()] for (W, b) in layer] for layer in params]
    flat = io.nested_to_flat(nested)
    assert flat.numel() == 144 and torch.equal(flat, O.flatten_layers(params))
    back = io.flat_to_nested(flat, 8, [1, 4, 2])
    for la, lb in zip(nested, back):
        for (Wa, ba), (Wb, bb) in zip(la, lb):
            assert np.array_equal(Wa, Wb) and np.array_equal(ba, bb)
    with pytest.raises(ValueError):
        io.flat_to_nested(flat[:-1], 8, [1, 4, 2])


def test_density_generators_shard_the_stream_by_rank(monkeypatch):
    """Host logic of densities.make_dataset_* (no GPU): batch b of a `world_size`-way run takes samples
    [b*batch + rank*local, +local) of the stream, so the ranks' slices tile the single-rank batch."""
    from normalizing_flows_b200 import densities, ops

    calls = []

    def fake(kind, seed, n, offset=0, device=None, **kw):
        calls.append((kind, seed, n, offset, kw))
        return torch.zeros(n, 2)

    monkeypatch.setattr(ops, "density_sample", fake)
    list(densities.make_dataset_energy_3(9, batch_size=96, num_batches=2))
    assert [(c[2], c[3]) for c in calls] == [(96, 0), (96, 96)]
    calls.clear()
    for rank in range(3):
        list(densities.make_dataset_gaussian_blobs(9, batch_size=96, num_batches=2, num_blobs=5, r=1.5, rank=rank, world_size=3))
    spans = sorted((c[3], c[3] + c[2]) for c in calls)
    assert spans == [(0, 32), (32, 64), (64, 96), (96, 128), (128, 160), (160, 192)]
    assert all(c[0] == "gaussian_blobs" and c[4] == {"num_blobs": 5, "r": 1.5} for c in calls)
    with pytest.raises(ValueError):
        list(densities.make_dataset_banana(1, batch_size=10, world_size=4))


def test_real_nvp_pack_params_cache(monkeypatch):
    """real_nvp.pack_params keeps the flat device buffer of a nested [[W1,b1],[W2,b2]] list while its leaves are
    unchanged (the reference passes the nested list on every forward_nvp / inverse_nvp call, real_nvp.py:71-134) and
    rebuilds it when a leaf is written in place or replaced.  CPU-only: the device copy is stubbed."""
    import numpy as np
    import torch

    from normalizing_flows_b200 import real_nvp as r

    copies = []
    monkeypatch.setattr(torch.Tensor, "to", lambda self, *a, **k: (copies.append(1) or self))
    r._PACK_CACHE.clear()
    p = r.init_mlp()
    f1 = r.pack_params([p])
    f2 = r.pack_params([p])  # a NEW outer list around the same leaves, as forward_nvp builds it
    assert f1 is f2 and len(copies) == 1
    p[0][0].add_(1.0)  # in-place write bumps the tensor's version
    f3 = r.pack_params([p])
    assert f3 is not f1 and abs(float(f3[0]) - (float(f1[0]) + 1.0)) < 1e-6
    p[1][1] = torch.ones(2)  # a replaced leaf
    f4 = r.pack_params([p])
    assert f4 is not f3 and float(f4[-1]) == 1.0
    q = [[np.ones((1, 4), np.float32), np.zeros(4, np.float32)], [np.ones((4, 2), np.float32), np.zeros(2, np.float32)]]
    g1 = r.pack_params([q])
    q[0][0][0, 0] = 5.0  # numpy leaves are fingerprinted by content
    g2 = r.pack_params([q])
    assert g2 is not g1 and float(g2[0]) == 5.0
    for _ in range(80):  # bounded
        r.pack_params([r.init_mlp()])
    assert len(r._PACK_CACHE) <= 64
