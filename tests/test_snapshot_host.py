"""var.dat snapshot writer/reader against an independent parser of gfortran's record framing and
against scipy.io.FortranFile (which reads what gfortran writes); host-only, no GPU."""
import os
import struct

import numpy as np
import pytest

from pencil_code_b200 import config as pcfg
from pencil_code_b200 import snapshot


def _records(path):
    """independent parser: list of record payloads; subrecords joined by the sign convention of gfortran
    (leading marker < 0: another subrecord follows; trailing marker < 0: one precedes)"""
    raw = open(path, "rb").read()
    recs, cur, pos, cont = [], b"", 0, False
    while pos < len(raw):
        (lead,) = struct.unpack_from("<i", raw, pos)
        n = abs(lead)
        payload = raw[pos + 4:pos + 4 + n]
        (trail,) = struct.unpack_from("<i", raw, pos + 4 + n)
        assert abs(trail) == n
        assert (trail < 0) == cont
        cur += payload
        pos += 8 + n
        cont = lead < 0
        if not cont:
            recs.append(cur)
            cur = b""
    assert not cont and pos == len(raw)
    return recs


def _case(n=(10, 6, 4), nvar=8):
    cfg = pcfg.make_config(n[0], True, True, ny=n[1], nz=n[2])
    rng = np.random.default_rng(7)
    f = rng.standard_normal((nvar, n[2] + 6, n[1] + 6, n[0] + 6))
    return cfg, f


def test_var_dat_layout(tmp_path):
    cfg, f = _case()
    path = snapshot.snapshot_path(str(tmp_path), 0)
    snapshot.output_snap(path, f, 1.25, cfg)
    recs = _records(path)
    assert len(recs) == 2
    assert np.array_equal(np.frombuffer(recs[0], dtype="<f8"), f.ravel())  # a(:,:,:,1:nvar), x fastest
    x, y, z = snapshot.mesh(cfg)
    tail = np.frombuffer(recs[1], dtype="<f8")
    assert np.array_equal(tail, np.concatenate([[1.25], x, y, z, [cfg["dx"], cfg["dy"], cfg["dz"]]]))
    assert x[3] == cfg["x0"] and np.allclose(np.diff(x), cfg["dx"])
    # and as gfortran-compatible readers see it
    from scipy.io import FortranFile
    with FortranFile(path, "r") as ff:
        a = ff.read_reals(np.float64)
        b = ff.read_reals(np.float64)
    assert np.array_equal(a, f.ravel()) and np.array_equal(b, tail)


def test_round_trip_and_overwrite(tmp_path):
    cfg, f = _case()
    path = str(tmp_path / "var.dat")
    snapshot.output_snap(path, f * 0, 0.0, cfg)
    snapshot.output_snap(path, f, 3.5, cfg)  # output_snap replaces an existing file
    g, t, (x, y, z), d3 = snapshot.input_snap(path, cfg)
    assert np.array_equal(g, f) and t == 3.5
    assert np.array_equal(x, snapshot.mesh(cfg)[0]) and d3 == (cfg["dx"], cfg["dy"], cfg["dz"])


@pytest.mark.parametrize("smax", [4096, 1000, 61440, 122880, 122872])
def test_subrecords(tmp_path, monkeypatch, smax):
    """records above the subrecord limit (2147483639 bytes; lowered through the test hook) are split and re-joined"""
    cfg, f = _case()
    monkeypatch.setenv("PCB200_SUBRECORD_MAX", str(smax))
    path = str(tmp_path / "var.dat")
    snapshot.output_snap(path, f, 2.0, cfg)
    nsub = -(-f.nbytes // smax)
    tail_bytes = 8 * (1 + 16 + 12 + 10 + 3)
    assert os.path.getsize(path) == f.nbytes + 8 * nsub + tail_bytes + 8 * -(-tail_bytes // smax)
    recs = _records(path)
    assert len(recs) == 2 and np.array_equal(np.frombuffer(recs[0], dtype="<f8"), f.ravel())
    g, t, _, _ = snapshot.input_snap(path, cfg)
    assert np.array_equal(g, f) and t == 2.0


def test_errors(tmp_path):
    cfg, f = _case()
    path = str(tmp_path / "var.dat")
    snapshot.output_snap(path, f, 0.0, cfg)
    with pytest.raises(IOError):
        snapshot.input_snap(str(tmp_path / "missing.dat"), cfg)
    with pytest.raises(IOError):  # wrong sizes: record structure does not match
        snapshot.input_snap(path, pcfg.make_config(12, True, True, ny=6, nz=4))
    with open(path, "r+b") as fh:  # truncated file
        fh.truncate(os.path.getsize(path) - 20)
    with pytest.raises(IOError):
        snapshot.input_snap(path, cfg)
    with pytest.raises(ValueError):
        snapshot.output_snap(path, f[:, :-1], 0.0, cfg)
