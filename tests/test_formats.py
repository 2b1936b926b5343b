"""Host-side file formats of the reference driver: byte-level checks."""
import numpy as np

from stanford_raytracer_b200 import rayfile


def test_es24_15e3():
    assert rayfile.es24(7.3712e6) == "  7.371200000000000E+006"
    assert rayfile.es24(-1.602e-19) == " -1.602000000000000E-019"
    assert rayfile.es24(0.0) == "  0.000000000000000E+000"
    assert rayfile.es24(1e-300) == "  1.000000000000000E-300"
    assert all(len(rayfile.es24(v)) == 24 for v in (1.0, -2.5e100, 3e-7, float("nan"), float("inf")))


def test_output_row_layout():
    row = np.arange(20, dtype=np.float64) + 0.5
    qs, ms, nus = [-1.602e-19] + [1.602e-19] * 3, [9.1e-31, 1.67e-27, 6.7e-27, 2.7e-26], [0.0] * 4
    line = rayfile.format_row(3, 1, row, 12566.370614359172, qs, ms, nus)
    # (2i10, 17es24.15e3, i10, 4*nspec es24.15e3) + newline = 822 + 1 characters for nspec = 4 (Appendix B.1)
    assert len(line) == 823 and line.endswith("\n")
    assert line[:20] == "         3         1"
    assert line[20:44] == rayfile.es24(0.5)
    assert line[20 + 16 * 24:20 + 17 * 24] == rayfile.es24(12566.370614359172)
    assert line[20 + 17 * 24:30 + 17 * 24] == "         4"
    body = line[30 + 17 * 24:-1]
    assert len(body) == 16 * 24
    assert body[8 * 24:9 * 24] == rayfile.es24(16.5)  # Ns(1) = row[16]


def test_read_inputrays(tmp_path):
    p = tmp_path / "rays.txt"
    p.write_text("1.0E007 0.0 0.0 0.0 0.0 1.0 6.28e4\n 7.3712d6, 0, 0, 0 0 0 12566.37 trailing-ignored\n\n5 5 5 1 0 0 100\nbad line stops reading\n9 9 9 0 0 1 7\n")
    pos0, dir0, w = rayfile.read_inputrays(str(p))
    assert pos0.shape == (3, 3)
    assert pos0[0, 0] == 1.0e7 and dir0[0, 2] == 1.0 and w[0] == 6.28e4
    assert pos0[1, 0] == 7.3712e6 and w[1] == 12566.37
