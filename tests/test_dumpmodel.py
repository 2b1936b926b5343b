"""`dumpmodel` (fortran/dumpmodel.f95): lattice construction and dump format on the CPU; on the GPU the dumped plasma
parameters against the oracle's funcPlasmaParams, bit for bit, through the file."""
import numpy as np
import pytest

import common
import oracle_lib as ol

R_E = 6371.2e3


def test_lattice_axis_is_the_references_linspace(built):
    from stanford_raytracer_b200 import dumpmodel as dm
    # (/ (ind, ind=0,nx-1) /)*delx + minx with delx = (maxx-minx)/(nx-1.0)   (dumpmodel.f95:219-226)
    x = dm.lattice_axis(7, -2.5 * R_E, 3.0 * R_E)
    d = (3.0 * R_E - -2.5 * R_E) / 6.0
    assert x.tolist() == [i * d + -2.5 * R_E for i in range(7)]
    assert dm.lattice_axis(1, 4.0, 9.0).tolist() == [4.0]  # a single node sits at min (:226)
    P = dm.lattice_points(3, 2, 2, (0.0, 2.0, 10.0, 11.0, 100.0, 101.0))
    assert P.shape == (12, 3)
    assert P[:4].tolist() == [[0.0, 10.0, 100.0], [1.0, 10.0, 100.0], [2.0, 10.0, 100.0], [0.0, 11.0, 100.0]]  # x fastest
    assert P[6].tolist() == [0.0, 10.0, 101.0]  # z slowest


def test_dump_file_format_round_trip(built, tmp_path):
    from stanford_raytracer_b200 import dumpmodel as dm
    rng = np.random.default_rng(3)
    f = rng.standard_normal((2, 3, 4, 19)) * 10.0 ** rng.integers(-300, 300, (2, 3, 4, 19))
    f[0, 0, 0, :4] = [-1.602e-19, 1.602e-19, 0.0, -0.0]
    ext = (-1.0 * R_E, 1.0 * R_E, 0.0, 0.0, -2.0 * R_E, 2.0 * R_E)
    p = tmp_path / "dump.txt"
    dm.write_dump(str(p), f, ext)
    lines = p.read_text().split("\n")
    assert lines[0] == "%10d%10d%10d%10d" % (4, 4, 3, 2)          # (5i10): nspec nx ny nz
    assert len(lines[1]) == 6 * 24 and lines[1][:24] == " -6.371200000000000E+006"  # (6es24.15e3)
    assert all(len(l) == 24 for l in lines[2:-1]) and lines[-1] == ""
    assert len(lines) - 3 == f.size                                # (es24.15e3): one value per record
    assert lines[2] == " -1.602000000000000E-019"
    back = dm.read_dump(str(p))
    assert (back["nspec"], back["nx"], back["ny"], back["nz"]) == (4, 4, 3, 2)
    assert np.array_equal(back["extent"], np.array(ext))
    # 16 significant digits do not round-trip every double; they do to 1 part in 1e15
    assert np.allclose(back["f"], f, rtol=1e-15, atol=0.0)
    again = tmp_path / "dump2.txt"
    dm.write_dump(str(again), back["f"], back["extent"])
    assert again.read_bytes() == p.read_bytes() or np.allclose(dm.read_dump(str(again))["f"], back["f"], rtol=1e-15, atol=0)


def test_dump_of_a_three_species_model_has_three_columns_per_quantity(built, tmp_path):
    """dumpmodel.f95:1284-1289 writes f(nspec*4+3, ...) with the model's own nspec: no phantom fourth species"""
    from stanford_raytracer_b200 import dumpmodel as dm

    class Fake:
        nspec = 3

        def funcPlasmaParams(self, x):
            raw = np.tile(np.arange(19.0), (len(x), 1))
            return {"raw": raw}

    f = dm.dump(Fake(), 2, 2, 1, (0.0, 1.0, 0.0, 1.0, 0.0, 0.0))
    assert f.shape == (1, 2, 2, 15)
    assert f[0, 0, 0].tolist() == [0, 1, 2, 4, 5, 6, 8, 9, 10, 12, 13, 14, 16, 17, 18]
    p = tmp_path / "d3.txt"
    dm.write_dump(str(p), f, (0.0, 1.0, 0.0, 1.0, 0.0, 0.0))
    assert p.read_text().split("\n")[0] == "%10d%10d%10d%10d" % (3, 2, 2, 1)
    assert dm.read_dump(str(p))["f"].shape == (1, 2, 2, 15)


@pytest.mark.gpu
def test_dumpmodel_matches_oracle_through_the_file(built, tmp_path):
    import stanford_raytracer_b200 as srt
    from stanford_raytracer_b200 import dumpmodel as dm
    srt.init(0)
    out = tmp_path / "ngo_dump.txt"
    ext = (-4.0 * R_E, 4.0 * R_E, 0.0, 0.0, -3.0 * R_E, 3.0 * R_E)
    rc = dm.main(["--modelnum", "1", "--ngo_configfile", common.NGO_CFG, "--yearday", "2010001", "--milliseconds_day", "0",
                  "--use_igrf", "0", "--use_tsyganenko", "0", "--minx", repr(ext[0]), "--maxx", repr(ext[1]),
                  "--minz", repr(ext[4]), "--maxz", repr(ext[5]), "--nx", "32", "--ny", "1", "--nz", "24", "--filename", str(out)])
    assert rc == 0
    d = dm.read_dump(str(out))
    assert (d["nx"], d["ny"], d["nz"], d["nspec"]) == (32, 1, 24, 4)
    P = dm.lattice_points(32, 1, 24, ext)
    keep = np.sqrt((P * P).sum(axis=1)) > 1.05 * R_E  # inside the Earth the analytic model is meaningless
    o = ol.OracleModel("ngo", common.NGO_CFG).plasma(np.ascontiguousarray(P[keep]))
    g = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0).funcPlasmaParams(P[keep])["raw"]
    assert np.array_equal(g, o, equal_nan=True)                                  # the kernel, bit for bit
    got = d["f"].reshape(-1, 19)[keep]
    fin = np.isfinite(o)
    assert np.array_equal(np.isfinite(got), fin)
    assert np.allclose(got[fin], o[fin], rtol=1e-15, atol=0.0)      # the file: 16 significant digits
    # and the text itself: every value is the ES24.15E3 rendering of the oracle's double
    from stanford_raytracer_b200.rayfile import es24
    lines = out.read_text().split("\n")[2:-1]
    idx = np.flatnonzero(keep)[:50]
    for n_, row in zip(idx, o[:50]):
        assert lines[19 * n_:19 * n_ + 19] == [es24(v) for v in row]
