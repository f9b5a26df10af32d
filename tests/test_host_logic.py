"""CPU tests (-m "not gpu") of the product's host side: the C-ABI library loads and exports every
symbol include/dcvc_b200.h declares, the host-built entropy tables equal the golden tables of the
unmodified reference, the scale->index thresholds reproduce build_indexes, the frame-file format
round-trips, and the module API keeps the reference's state_dict keys.  No kernel is launched."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

import dcvc_b200
from dcvc_b200 import _lib, arch, entropy_models as em, stream_helper as sh
from dcvc_b200.common_model import get_downsampled_shape, get_padding_size

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "dcvc_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(dcvc_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    names = _declared_symbols()
    assert len(names) >= 30
    so = ctypes.CDLL(os.path.join(ROOT, "dcvc_b200", "libdcvc_b200.so"))
    missing = [n for n in names if not hasattr(so, n)]
    assert not missing, missing
    unbound = [n for n in names if n not in _lib.SIGNATURES]
    assert not unbound, f"declared in the header but not bound by _lib.SIGNATURES: {unbound}"
    undeclared = [n for n in _lib.SIGNATURES if n not in names]
    assert not undeclared, f"bound but not declared in the header: {undeclared}"


def test_version_and_error_string():
    lib = _lib.lib()
    assert lib.dcvc_version() >= 1
    # a bad-argument call reports through the status code + dcvc_last_error, never throws
    rc = lib.dcvc_pmf_to_quantized_cdf(None, 0, 16, None)
    assert rc != 0
    assert len(lib.dcvc_last_error()) > 0
    with pytest.raises(RuntimeError):
        _lib.check(rc)


def test_no_cpu_fallback():
    """Without a CUDA device the product refuses to code (it must not route to the oracle)."""
    if torch.cuda.is_available():
        pytest.skip("CPU-only check")
    net = dcvc_b200.IntraNoAR()
    net.load_state_dict(arch.synth_state_dict(arch.intra_spec(), 0))
    with pytest.raises(RuntimeError):
        net.update(force=True)
    src = open(os.path.join(ROOT, "dcvc_b200", "__init__.py")).read()
    for f in os.listdir(os.path.join(ROOT, "dcvc_b200")):
        if f.endswith(".py"):
            src += open(os.path.join(ROOT, "dcvc_b200", f)).read()
    assert "import oracle" not in src and "from oracle" not in src


def test_pmf_to_quantized_cdf_properties():
    rng = np.random.default_rng(0)
    for n in (2, 3, 17, 101):
        pmf = rng.random(n).astype(np.float32) ** 4
        pmf /= pmf.sum()
        cdf = em.pmf_to_quantized_cdf(pmf)
        assert cdf[0] == 0 and cdf[-1] == 1 << 16
        assert np.all(np.diff(cdf.astype(np.int64)) > 0)       # every symbol stays codable


def test_gaussian_tables_equal_reference_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "dc_tables.npz"))
    for dist in ("laplace", "gaussian"):
        cdf, sizes, offs = em.gaussian_tables(dist)
        assert np.array_equal(cdf, g[dist + "_cdf"])
        assert np.array_equal(sizes, g[dist + "_len"])
        assert np.array_equal(offs, g[dist + "_off"])


def _bitparm(sd, prefix):
    out = {}
    for k in ("f1", "f2", "f3", "f4"):
        names = ("h", "b", "a") if k != "f4" else ("h", "b")
        out[k] = tuple(sd[f"{prefix}.{k}.{n}"].float() for n in names)
    return out


def test_factorized_tables_equal_reference_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "dc_tables.npz"))
    sd_i = arch.synth_state_dict(arch.intra_spec(), 0)
    sd_p = arch.synth_state_dict(arch.dmc_spec(), 100)
    for tag, sd, p in (("i_z", sd_i, "bit_estimator_z"), ("p_z", sd_p, "bit_estimator_z"),
                       ("p_z_mv", sd_p, "bit_estimator_z_mv")):
        cdf, sizes, offs = em.factorized_tables(_bitparm(sd, p))
        assert np.array_equal(cdf, g[tag + "_cdf"]), tag
        assert np.array_equal(sizes, g[tag + "_len"]), tag
        assert np.array_equal(offs, g[tag + "_off"]), tag


@pytest.mark.parametrize("dist", ["laplace", "gaussian"])
def test_index_edges_reproduce_build_indexes(dist):
    """index = (#edges <= scale) - 1 must equal GaussianEncoder.build_indexes for any fp32 scale:
    random scales, every edge and both of its fp32 neighbours."""
    edges = em.index_edges(dist)
    assert edges.shape == (256,) and float(edges[0]) == 0.0
    assert bool(torch.all(edges[1:] > edges[:-1]))
    g = torch.Generator().manual_seed(3)
    s = torch.exp(torch.rand(200000, generator=g) * 12.0 - 6.0)
    bits = edges[1:].view(torch.int32)
    near = torch.cat([(bits + d).view(torch.float32) for d in (-2, -1, 0, 1, 2)])
    s = torch.cat([s, near, torch.tensor([0.0, 1e-9, 1e-5, 0.01, 0.11, 64.0, 1e4])])
    want = em._index_of(s, dist)
    got = (torch.searchsorted(edges, s, right=True) - 1).int()
    assert torch.equal(got, want)


def test_fine_q_tables_equal_reference_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "dc_tables.npz"))
    i_net = dcvc_b200.IntraNoAR()
    i_net.load_state_dict(arch.synth_state_dict(arch.intra_spec(), 0))
    p_net = dcvc_b200.DMC()
    p_net.load_state_dict(arch.synth_state_dict(arch.dmc_spec(), 100))
    i_net._build_fine_tables()
    p_net._build_fine_tables()
    assert np.array_equal(i_net._fine["q_scale_enc"], g["i_q_enc_fine"])
    assert np.array_equal(i_net._fine["q_scale_dec"], g["i_q_dec_fine"])
    assert np.array_equal(p_net._fine["y_q_scale_enc"], g["p_y_q_enc_fine"])
    assert np.array_equal(p_net._fine["mv_y_q_scale_dec"], g["p_mv_y_q_dec_fine"])


def test_state_dict_keys_match_oracle():
    """A reference checkpoint must load unchanged: same parameter names and shapes as the
    architecture spec the oracle (pinned to the reference) consumes."""
    for cls, spec in ((dcvc_b200.IntraNoAR, arch.intra_spec()), (dcvc_b200.DMC, arch.dmc_spec())):
        sd = arch.synth_state_dict(spec, 7)
        net = cls()
        own = net.state_dict()
        assert set(own.keys()) == set(sd.keys())
        for k, v in sd.items():
            assert tuple(own[k].shape) == tuple(v.shape), k
        net.load_state_dict(sd)
        bad = dict(sd)
        bad.pop(next(iter(bad)))
        with pytest.raises((RuntimeError, KeyError)):
            net.load_state_dict(bad)


def test_frame_file_format_round_trip(tmp_path):
    payload = bytes(range(256)) * 3
    blob = sh.pack_i(1080, 1920, True, 3, payload)
    assert len(blob) == 13 + len(payload)
    assert blob[:8] == (1080).to_bytes(4, "big") + (1920).to_bytes(4, "big")
    assert blob[8] == (1 << 7) | (3 << 1)
    assert sh.unpack_i(blob) == (1080, 1920, True, 3, payload)
    blob = sh.pack_p(payload, False, 63, 2)
    assert len(blob) == 6 + len(payload) and blob[0] == 63 << 1 and blob[1] == 2
    assert sh.unpack_p(blob) == (False, 63, 2, payload)
    with pytest.raises(ValueError):
        sh.unpack_p(blob[:-1])
    path = str(tmp_path / "0.bin")
    sh.encode_i(64, 96, False, 17, payload, path)
    assert sh.filesize(path) == 13 + len(payload)
    assert sh.decode_i(path) == (64, 96, False, 17, payload)
    sh.encode_p(b"", True, 0, 1, path)                     # empty payload is legal
    assert sh.decode_p(path) == (True, 0, 1, b"")
    with pytest.raises(ValueError):
        sh.filesize(str(tmp_path / "missing.bin"))


def test_padding_and_latent_shapes():
    assert get_padding_size(1080, 1920, 16) == (0, 0, 0, 8)
    assert get_padding_size(1088, 1920, 64) == (0, 0, 0, 0)
    assert get_downsampled_shape(1080, 1920, 64) == (17, 30)
    assert get_downsampled_shape(2160, 3840, 16) == (135, 240)


def test_search_lut_brackets_every_cumulative_value():
    """dcvc_tables.lut (entropy_models.search_lut): for every 16-bit cumulative value the symbol
    the reference's linear search finds (rans.cpp:289-292) lies in [lut[b], lut[b + 1]] of its
    bucket b = cum >> 8 (the row's last symbol closing the last bucket), on the real Gaussian /
    Laplace tables and on a row of all-ones-but-one frequencies."""
    import numpy as np
    from dcvc_b200 import entropy_models as em
    for dist in ("gaussian", "laplace"):
        cdf, sizes, _ = em.gaussian_tables(dist)
        lut = em.search_lut(cdf, sizes)
        assert lut is not None and lut.shape == (cdf.shape[0], 256)
        cum = np.arange(65536, dtype=np.int64)
        for r in (0, 1, cdf.shape[0] // 2, cdf.shape[0] - 1):
            n = int(sizes[r])
            row = np.asarray(cdf[r, :n], dtype=np.int64)
            assert row[0] == 0 and row[n - 1] == 65536
            sym = np.searchsorted(row, cum, side="right") - 1          # largest s with cdf[s] <= cum
            lo = lut[r][cum >> 8].astype(np.int64)
            hi = np.where((cum >> 8) < 255, lut[r][np.minimum((cum >> 8) + 1, 255)], n - 2)
            assert np.all(lo <= sym) and np.all(sym <= hi), (dist, r)
    # many one-count symbols in one bucket
    n = 200
    row = np.concatenate(([0], np.cumsum(np.r_[np.ones(n - 2, dtype=np.int64), 65536 - (n - 2)])))
    cdf = row[None, :].astype(np.int32)
    lut = em.search_lut(cdf, np.array([n], dtype=np.int32))
    assert lut[0, 0] == 0 and lut[0, 1] == n - 2 and lut[0, 255] == n - 2
    assert em.search_lut(np.zeros((1, 300), dtype=np.int32), np.array([300])) is None


def test_graph_helpers_recognise_dpb_views_and_placement():
    """dcvc_b200.graphs: a dpb entry (an NCHW view of an NHWC buffer, possibly a channel slice) is
    recognised with its buffer, channel offset and count; tensors that are not such views are
    not; tensor.placed hands out the preassigned buffer only for the layout asked for."""
    import torch
    from dcvc_b200 import graphs, tensor
    from dcvc_b200.tensor import Act
    buf = torch.arange(6 * 10 * 8, dtype=torch.float32).reshape(6, 10, 8)
    a = Act(buf, 4, 3)                                   # channels 4..6 of an 8-wide buffer
    v = a.nchw()
    assert tuple(v.shape) == (1, 3, 6, 10)
    back = graphs._as_act(v)
    assert back is not None and back.buf.data_ptr() == buf.data_ptr()
    assert (back.c0, back.c, back.h, back.w) == (4, 3, 6, 10)
    assert torch.equal(back.buf, buf)
    assert graphs._as_act(torch.zeros(1, 3, 6, 10)) is None            # plain NCHW: not a view
    assert graphs._as_act(torch.zeros(2, 3, 6, 10)) is None
    s = graphs._Static({"ref_frame": v, "ref_y": None})
    assert s.views["ref_y"] is None and tuple(s.views["ref_frame"].shape) == (1, 3, 6, 10)
    assert s.bufs["ref_frame"][0].shape == buf.shape and s.acts["ref_frame"].c0 == 4
    s.load({"ref_frame": v, "ref_y": None})
    assert torch.equal(s.views["ref_frame"], v)
    dpb = {k: None for k in graphs.DPB_KEYS}
    dpb["ref_frame"] = s.views["ref_frame"]
    assert graphs._is_static(dpb)
    dpb["ref_frame"] = v
    assert not graphs._is_static(dpb)
    assert graphs._sig(dpb) != graphs._sig({**dpb, "ref_frame": s.views["ref_frame"]})
    assert tensor.placed("ref_y", 6, 10, 3) is None
    tensor.PLACEMENT = {"ref_y": Act(buf, 0, 8)}
    try:
        assert tensor.placed("ref_y", 6, 10, 8).buf.data_ptr() == buf.data_ptr()
        assert tensor.placed("other", 6, 10, 8) is None
        try:
            tensor.placed("ref_y", 6, 10, 4)
            raise RuntimeError("a layout mismatch must be refused")
        except AssertionError:
            pass
    finally:
        tensor.PLACEMENT = None
