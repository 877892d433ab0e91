"""`--onnx_path` loading (fangyan_tts_b200/weights.py): an exported graph carries neither our tensor names nor our
layouts, so .onnx files are mapped by topology.  The graphs come from tests/onnx_writer.py (exporter conventions:
MatMul weights [in, out] under `onnx::MatMul_N`, BatchNorm folded into the preceding Conv, other projects' names)."""
import numpy as np
import pytest
import torch

from fangyan_tts_b200 import synth_weights, weights
from fangyan_tts_b200.configs import CamPlusConfig, S3Config
from oracle import campplus as CP

import onnx_writer as OW

SMALL_S3 = dict(d_model=128, n_head=2, d_ff=256, n_layer=2)


def test_s3tok_onnx_maps_by_topology_bit_exact(tmp_path):
    cfg = S3Config(**SMALL_S3)
    st = synth_weights.s3tok_state(cfg, seed=3)
    OW.write_s3tok(tmp_path / "speech_tokenizer_v3.onnx", st, cfg)
    raw = weights.read_onnx_initializers(tmp_path / "speech_tokenizer_v3.onnx")
    assert not any(k in raw for k in st), "the file must not carry our names (the point of the test)"
    assert any(k.startswith("onnx::MatMul_") for k in raw)
    got = weights.load_state(tmp_path / "speech_tokenizer_v3.onnx", "s3tok", cfg)
    assert sorted(got) == sorted(st)
    for k in st:                                       # un-transposed, unrenamed, nothing lost
        assert got[k].shape == st[k].shape and np.array_equal(got[k], st[k]), k
    weights.check_shapes(got, "s3tok", cfg)


def test_s3tok_onnx_structure_errors_are_structural(tmp_path):
    cfg = S3Config(**SMALL_S3)
    st = synth_weights.s3tok_state(cfg, seed=3)
    OW.write_s3tok(tmp_path / "a.onnx", st, cfg)
    with pytest.raises(weights.OnnxStructureError, match=r"blocks\.2\.attn_ln"):        # table says 3 layers, graph has 2
        weights.load_state(tmp_path / "a.onnx", "s3tok", S3Config(**{**SMALL_S3, "n_layer": 3}))
    with pytest.raises(weights.OnnxStructureError, match=r"fc1.*Linear\[512, 128\]"):  # table says d_ff 512
        weights.load_state(tmp_path / "a.onnx", "s3tok", S3Config(**{**SMALL_S3, "d_ff": 512}))
    OW.write_s3tok(tmp_path / "b.onnx", st, cfg, swap_kv=True)                          # value before key: bias pattern breaks
    with pytest.raises(weights.OnnxStructureError, match=r"attn\.key"):
        weights.load_state(tmp_path / "b.onnx", "s3tok", cfg)
    OW.write_s3tok(tmp_path / "c.onnx", st, cfg, ln_eps=1e-6)                           # ADVICE: eps of the real model unknown
    with pytest.raises(weights.OnnxStructureError, match="epsilon 1e-06"):
        weights.load_state(tmp_path / "c.onnx", "s3tok", cfg)
    got = weights.load_state(tmp_path / "c.onnx", "s3tok", S3Config(**{**SMALL_S3, "ln_eps": 1e-6}))
    assert np.array_equal(got["blocks.1.fc2.weight"], st["blocks.1.fc2.weight"])
    # the two norms of a block may carry different epsilons: the table has one field for each
    OW.write_s3tok(tmp_path / "d.onnx", st, cfg, ln_eps=1e-6, mlp_ln_eps=1e-5)
    with pytest.raises(weights.OnnxStructureError, match=r"mlp_ln has epsilon 1e-05"):
        weights.load_state(tmp_path / "d.onnx", "s3tok", S3Config(**{**SMALL_S3, "ln_eps": 1e-6}))
    weights.load_state(tmp_path / "d.onnx", "s3tok", S3Config(**{**SMALL_S3, "ln_eps": 1e-6, "mlp_ln_eps": 1e-5}))


def _cam(state):
    m = CP.CamPlus()
    m.load_state_dict({k: torch.from_numpy(np.ascontiguousarray(v)) for k, v in state.items()}, strict=False)
    return m.eval()


@pytest.mark.parametrize("fold,gemm", [(True, False), (False, False), (True, True)])
def test_campplus_onnx_folded_and_unfolded_compute_the_same_network(tmp_path, fold, gemm):
    cfg = CamPlusConfig()
    st = synth_weights.campplus_state(cfg, seed=5)
    OW.write_campplus(tmp_path / "campplus.onnx", st, cfg, fold=fold, dense_as_gemm=gemm)
    raw = weights.read_onnx_initializers(tmp_path / "campplus.onnx")
    assert not any(k in raw for k in st if k.startswith("blocks."))
    got = weights.load_state(tmp_path / "campplus.onnx", "campplus", cfg)
    weights.check_shapes(got, "campplus", cfg)
    if not fold:
        for k in st:
            assert np.allclose(got[k], st[k], rtol=0, atol=1e-7), k
    feat = torch.from_numpy(np.random.default_rng(1).standard_normal((1, 150, 80)).astype(np.float32))
    with torch.no_grad():
        a, b = _cam(st)(feat), _cam(got)(feat)
    assert float(torch.nn.functional.cosine_similarity(a, b).min()) > 0.999999
    assert float((a - b).abs().max()) <= 2e-4 * float(a.abs().max())


def test_campplus_onnx_missing_layer_is_a_structural_error(tmp_path):
    cfg = CamPlusConfig()
    st = synth_weights.campplus_state(cfg, seed=5)
    OW.write_campplus(tmp_path / "x.onnx", st, cfg, drop_layer=(1, 7))
    with pytest.raises(weights.OnnxStructureError, match=r"blocks\.1\.7"):
        weights.load_state(tmp_path / "x.onnx", "campplus", cfg)


def test_check_shapes_catches_transposed_square_and_missing():
    cfg = S3Config(**SMALL_S3)
    st = synth_weights.s3tok_state(cfg, seed=3)
    weights.check_shapes(st, "s3tok", cfg)
    bad = dict(st)
    bad["conv1.weight"] = np.ascontiguousarray(st["conv1.weight"].transpose(0, 2, 1))     # same numel, wrong layout
    with pytest.raises(ValueError, match="conv1.weight"):
        weights.check_shapes(bad, "s3tok", cfg)
    del bad["conv1.weight"]
    with pytest.raises(KeyError, match="conv1.weight"):
        weights.check_shapes(bad, "s3tok", cfg)
