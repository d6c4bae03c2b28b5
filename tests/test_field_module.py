"""CPU: host logic of the NeuralDisplacementField drop-in (find_b200/field.py) — constructor determinism against
the reference's own class (hash recorded in tests/golden/field_golden.npz by oracle/make_golden.py), state_dict keys
and shapes, LatentVector indexing (model.py:130-152), loud failure without CUDA."""
import hashlib

import numpy as np
import pytest
import torch

from find_b200.field import LatentVector, NeuralDisplacementField, sample_latent_vectors


def _make(**kw):
    args = dict(use_shapevec=True, use_texvec=True, use_posevec=True, train_size=8, val_size=2, shapevec_size=100,
                texvec_size=100, posevec_size=100, device="cpu", opts=None)
    args.update(kw)
    return NeuralDisplacementField(**args)


def _state_hash(sd):
    h = hashlib.sha256()
    for k in sorted(sd):
        h.update(k.encode())
        h.update(sd[k].detach().cpu().contiguous().numpy().tobytes())
    return h.hexdigest()


def test_ctor_state_is_bit_identical_to_reference(golden):
    m = _make()
    assert _state_hash(m.state_dict()) == str(golden["ctor_state_sha256"])
    np.testing.assert_array_equal(m.encoder[0]._B.numpy(), golden["B"])
    np.testing.assert_allclose(m.base[0].weight[0, :3].detach().numpy(), [-0.0071, 0.0176, -0.0338], atol=1e-4)


def test_state_dict_keys_and_shapes():
    sd = _make().state_dict()
    expect = {"template_verts": (1, 1, 3), "template_faces": (1, 1, 3), "avg_col": (3,),
              "base.0.weight": (256, 515), "base.8.bias": (256,), "mlp_disp.0.weight": (256, 456),
              "mlp_col.0.weight": (256, 356), "mlp_disp.6.weight": (3, 256), "mlp_col.6.bias": (3,),
              "shapevec.data": (8, 100), "posevec_val.data": (2, 100), "texvec.data": (8, 100), "reg.data": (8, 9),
              "reg_val.data": (2, 9)}
    for k, shp in expect.items():
        assert tuple(sd[k].shape) == shp, k
    assert not any(k.startswith("encoder") for k in sd)              # _B is not in the state_dict (reference quirk)
    assert sum(v.numel() for k, v in sd.items() if k.startswith(("base.", "mlp_"))) == 868358
    assert torch.equal(sd["reg.data"][0], torch.tensor([0, 0, 0, 0, 0, 0, 1, 1, 1.0]))
    assert float(sd["mlp_disp.6.weight"].abs().max()) == 0.0         # reset_weights, model.py:516-518


def test_param_groups_match_reference_lists():
    m = _make()
    assert len(m.main_params) == 26 + 3
    assert len(m.latent_params) == 6 and len(m.reg_params) == 2 and len(m.val_params) == 3
    assert [v.name for v in m.latent_vectors_train] == ["shapevec_train", "posevec_train", "texvec_train", "reg_train"]
    assert len(m.mlp_parameters()) == 26


def test_latent_vector_indexing_and_sampling():
    lv = LatentVector(vec_size=4, name="shapevec_train", device="cpu", key="shape", labels=["a", "b", "a2", "b"])
    with torch.no_grad():
        lv.data.copy_(torch.arange(16.0).reshape(4, 4))
    assert torch.equal(lv["b"], lv.data[1])                       # list.index semantics: first occurrence
    assert torch.equal(lv[["a2", "a"]], lv.data[[2, 0]])
    assert torch.equal(lv[torch.tensor([3, 3])], lv.data[[3, 3]])
    assert torch.equal(lv[2], lv.data[2])
    with pytest.raises(ValueError):
        lv["zzz"]
    plain = LatentVector(5, vec_size=3, name="reg_train", device="cpu", key="reg")
    out = sample_latent_vectors({"idx": torch.tensor([4, 1]), "shape": ["b", "a"]}, [lv, plain])
    assert torch.equal(out["shapevec_train"], lv.data[[1, 0]]) and out["reg_train"].shape == (2, 3)
    with pytest.raises(AssertionError):
        plain["a"]


def test_unsupported_architectures_fail_loudly():
    for kw in (dict(width=128), dict(depth=3), dict(dispdepth=2), dict(progressive_encoding=True),
               dict(positional_encoding=False), dict(input_dim=2)):
        with pytest.raises(NotImplementedError):
            _make(**kw)


def test_no_cpu_path():
    m = _make()
    with pytest.raises(RuntimeError, match="CUDA"):
        m(torch.zeros(1, 4, 3), shapevec=torch.zeros(2, 100), texvec=torch.zeros(2, 100), posevec=torch.zeros(2, 100))


def test_save_load_roundtrip(tmp_path):
    m = _make()
    m.set_template(torch.randn(7, 3), torch.zeros(2, 3, dtype=torch.int64))
    with torch.no_grad():
        m.shapevec.data.normal_()
    m.save_model(str(tmp_path), "ck")

    class O:
        dont_load_latents = False
        template_features_pth = None
    m2 = NeuralDisplacementField.load(str(tmp_path / "ck.pth"), device="cpu", opts=O())
    for (k, a), (_, b) in zip(m.state_dict().items(), m2.state_dict().items()):
        assert torch.equal(a, b), k
