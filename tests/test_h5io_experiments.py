"""§8(f4) callers & formats: the HDF5 files the reference's scripts exchange (gpt_b200/h5io.py) and the experiment driver
(gpt_b200/experiments.py)."""
import os

import numpy as np
import pytest

import h5lite
from gpt_b200 import h5io
from util import load_golden


def test_h5_round_trip_through_two_readers(tmp_path):
    """write_h5 output is read back identically by the product's reader and by the oracle's independent reader (the one
    validated against every .h5 file of the reference); more datasets than one symbol-table node holds; Julia
    orientation of multi-dimensional arrays."""
    rng = np.random.default_rng(0)
    d = {"testRMSE": rng.random(200), "w_store": rng.random((20, 7)), "U_store": np.asfortranarray(rng.random((12, 3, 4, 7))),
         "I": rng.integers(1, 5, (20, 4)).astype(np.int32), "f32": rng.random((3, 2)).astype(np.float32), "i64": np.arange(5)}
    for i in range(11):
        d[f"extra{i}"] = rng.random(i + 1)
    p = str(tmp_path / "x.h5")
    h5io.write_h5(p, d)
    a, b = h5io.read_h5(p), h5lite.read_all(p)
    assert sorted(a) == sorted(b) == sorted(d)
    for k, v in d.items():
        assert np.array_equal(a[k], v) and np.array_equal(b[k], v) and a[k].dtype == np.asarray(v).dtype, k
    # element [j,l,k,t] as Julia indexes U_store sits at the column-major offset in the file
    raw = open(p, "rb").read()
    U = d["U_store"]
    off = raw.index(U.ravel(order="F")[:4].tobytes())
    assert np.frombuffer(raw, "<f8", count=U.size, offset=off).reshape(U.shape, order="F")[5, 2, 3, 6] == U[5, 2, 3, 6]
    with pytest.raises(ValueError):
        h5io.write_h5(p, {"a/b": np.zeros(3)})


def test_h5_messages_match_the_reference_file(tmp_path):
    """The object header of a Float64 vector carries the same dataspace / datatype / fill-value / layout messages as the
    reference's own testRMSE_PP.h5 (bytes copied from that file; the layout message's address differs by construction)."""
    p = str(tmp_path / "t.h5")
    h5io.write_h5(p, {"testRMSE": np.arange(200, dtype=np.float64)})
    raw = open(p, "rb").read()
    assert raw[:8] == b"\x89HDF\r\n\x1a\n" and raw[8:24].hex() == "00000000000808000400100000000000"   # superblock bytes 8..23
    want = {
        "dataspace": "0100180000000000" "0101010000000000c800000000000000c800000000000000",
        "datatype": "0300180001000000" "11203f000800000000004000340b0034ff03000000000000",
        "fill": "0500080001000000" "0202020100000000",
    }
    for k, hx in want.items():
        assert bytes.fromhex(hx) in raw, k
    i = raw.index(bytes.fromhex("0800180000000000" "0301"))     # contiguous layout v3: address, then 200*8 bytes
    assert int.from_bytes(raw[i + 18:i + 26], "little") == 1600
    hp = raw.index(b"HEAP")
    free_off, seg = int.from_bytes(raw[hp + 16:hp + 24], "little"), int.from_bytes(raw[hp + 24:hp + 32], "little")
    assert int.from_bytes(raw[seg + free_off:seg + free_off + 8], "little") == 1      # free list ends (H5HL_FREE_NULL)


@pytest.mark.gpu
@pytest.mark.parametrize("sampler", ["sgld", "rmsprop"])
def test_powerplant_experiment_against_the_stored_curves(sampler, tmp_path):
    """The PowerPlant experiment end to end (whiten -> feature -> sampler -> per-epoch pred -> RMSE curve -> h5) on the
    reference's data, against the reference's own stored 200-epoch curves (testRMSE_PP.h5: `testRMSE` for SGLD,
    `testRMSE2` for RMSprop; tests/golden/rmse_curves.npz).  Different random draws (Julia's RNG cannot be reproduced),
    so the comparison is statistical: mean of the last 50 epochs within 4 %, best epoch within 4 %, and the same
    shape — a clear descent from the first epochs."""
    from gpt_b200 import experiments as E
    pp = load_golden("powerplant.npz")
    ref = load_golden("rmse_curves.npz")["pp_sgld" if sampler == "sgld" else "pp_rmsprop"]
    # the fixture holds the reference's PPdata_full.h5: inputs and targets already standardised (y / 17.067); seed 1 is
    # a start that converges (the reference's step sizes sit at the stability limit: see regression_experiment)
    out = E.powerplant_experiment(pp["Xtrain"], pp["ytrain"], sampler=sampler, seed=1, target_scale=17.06699,
                                  out_prefix=str(tmp_path / "pp_"))
    curve = out["testRMSE"]
    assert curve.shape == (200,) and np.isfinite(curve).all()
    assert abs(curve[-50:].mean() / ref[-50:].mean() - 1) < 0.04, (curve[-50:].mean(), ref[-50:].mean())
    assert abs(curve.min() / ref.min() - 1) < 0.04, (curve.min(), ref.min())
    assert curve[:3].mean() > 1.05 * curve[-50:].mean()
    assert out["final_rmse"] <= curve[-50:].mean() * 1.001            # the averaged prediction beats the average error
    back = h5io.read_h5(str(tmp_path / "pp_testRMSE.h5"))
    assert np.array_equal(back["testRMSE"], curve)
    for k in range(4):
        Uk = out["U_store"][:, :, k, -1]
        assert np.abs(Uk.T @ Uk - np.eye(20)).max() < 1e-9
