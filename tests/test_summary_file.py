"""CPU: summary.shopt writer / reader (dataOutprocessing.jl:52-84, reader.jl:13-38).  The file the library writes is
parsed by an independent minimal FITS parser (below), and a file in the reference's FULL layout, assembled here in
Python, is read back by the library.  Host code: no device needed."""
import struct

import numpy as np


def parse_fits(path):
    """[(header dict, raw data bytes)] -- 2880-byte blocks, 80-character cards, END."""
    raw = open(path, "rb").read()
    assert len(raw) % 2880 == 0
    off, hdus = 0, []
    while off < len(raw):
        hdr, done = {}, False
        while not done:
            block = raw[off:off + 2880].decode("ascii")
            off += 2880
            for i in range(36):
                card = block[80 * i:80 * i + 80]
                key = card[:8].strip()
                if key == "END":
                    done = True
                    break
                if card[8:10] == "= ":
                    val = card[10:].split("/")[0].strip() if "'" not in card[10:] else card[10:].split("'")[1].strip()
                    hdr[key] = val
        naxis = int(hdr["NAXIS"])
        n = 1 if naxis else 0
        for a in range(1, naxis + 1):
            n *= int(hdr["NAXIS%d" % a])
        nbytes = abs(int(hdr["BITPIX"])) // 8 * int(hdr.get("GCOUNT", 1)) * (int(hdr.get("PCOUNT", 0)) + n)
        hdus.append((hdr, raw[off:off + nbytes]))
        off += (nbytes + 2879) // 2880 * 2880
    return hdus


def card(key, value, quoted=False):
    s = "%-8s= '%-8s'" % (key, value) if quoted else "%-8s= %20s" % (key, value)
    return s.ljust(80)


def hdu_bytes(cards, data):
    h = "".join(cards) + "END".ljust(80)
    h = h.ljust((len(h) + 2879) // 2880 * 2880)
    return h.encode("ascii") + data + b"\0" * (-len(data) % 2880)


def table(extname, cols):
    rows = max(len(v) for _, v in cols)
    cards = [card("XTENSION", "BINTABLE", True), card("BITPIX", 8), card("NAXIS", 2), card("NAXIS1", 8 * len(cols)),
             card("NAXIS2", rows), card("PCOUNT", 0), card("GCOUNT", 1), card("TFIELDS", len(cols))]
    for i, (name, _) in enumerate(cols):
        cards += [card("TTYPE%d" % (i + 1), name, True), card("TFORM%d" % (i + 1), "D", True)]
    cards.append(card("EXTNAME", extname, True))
    data = b"".join(struct.pack(">d", float(v[r]) if r < len(v) else 0.0) for r in range(rows) for _, v in cols)
    return hdu_bytes(cards, data)


def image(extname, arr, primary=False):
    arr = np.ascontiguousarray(arr, dtype=">f8")
    cards = [card("SIMPLE", "T") if primary else card("XTENSION", "IMAGE", True), card("BITPIX", -64),
             card("NAXIS", arr.ndim)]
    cards += [card("NAXIS%d" % (i + 1), n) for i, n in enumerate(arr.shape[::-1])]
    cards += [card("EXTEND", "T")] if primary else [card("PCOUNT", 0), card("GCOUNT", 1)]
    cards.append(card("EXTNAME", extname, True))
    return hdu_bytes(cards, arr.tobytes())


def test_written_file_has_the_reference_layout(stage, tmp_path):
    rng = np.random.default_rng(0)
    degree, r, c = 3, 7, 5
    K = 10
    poly = rng.normal(size=(K, c, r))
    field = rng.normal(size=(3, 10))
    path = tmp_path / "summary.shopt"
    stage.write_summary(path, poly, degree, field)
    hdus = parse_fits(path)
    assert [h["EXTNAME"] for h, _ in hdus] == ["POLYNOMIAL_MATRIX", "SUMMARY_STATISTICS", "ANALYTIC_PROFILE"]
    h0, d0 = hdus[0]
    assert (h0["SIMPLE"], h0["BITPIX"], h0["NAXIS"]) == ("T", "-64", "3")
    assert (int(h0["NAXIS1"]), int(h0["NAXIS2"]), int(h0["NAXIS3"])) == (K, c, r)      # numpy shape (r, c, K)
    cube = np.frombuffer(d0, dtype=">f8").reshape(r, c, K)                                 # what astropy hands to Julia
    assert np.array_equal(cube, poly.transpose(2, 1, 0))                                   # PolynomialMatrix[a, b, k]
    h1, d1 = hdus[1]
    assert h1["TTYPE1"] == "POLYNOMIAL_DEGREE" and struct.unpack(">d", d1[:8])[0] == 3.0
    h2, d2 = hdus[2]
    assert [h2["TTYPE%d" % i] for i in (1, 2, 3)] == ["s_MATRIX", "g1_MATRIX", "g2_MATRIX"] and int(h2["NAXIS2"]) == 10
    assert np.array_equal(np.frombuffer(d2, dtype=">f8").reshape(10, 3).T, field)
    got_poly, got_deg, got_field = stage.read_summary(path)
    assert got_deg == 3 and np.array_equal(got_poly, poly) and np.array_equal(got_field, field)


def test_reads_the_full_untruncated_layout(stage, tmp_path):
    """dataOutprocessing.jl:86-168: seven HDUs, ANALYTIC_PROFILE last, POLYNOMIAL_DEGREE the 13th column of a table
    whose other columns have one row per star."""
    rng = np.random.default_rng(1)
    degree, r, c, n = 2, 6, 6, 23
    K = 6
    cube = rng.normal(size=(r, c, K))
    field = rng.normal(size=(3, 10))
    names = ["U_COORDINATES", "V_COORDINATES", "TRAINING_U_COORDS", "TRAINING_V_COORDS", "VALIDATION_U_COORDS",
             "VALIDATION_V_COORDS", "S_MODEL", "g1_MODEL", "g2_MODEL", "s_DATA", "g1_DATA", "g2_DATA"]
    cols = [(nm, rng.normal(size=n)) for nm in names] + [("POLYNOMIAL_DEGREE", [float(degree)])]
    blob = (image("POLYNOMIAL_MATRIX", cube, primary=True) + table("SUMMARY_STATISTICS", cols)
            + image("VIGNETS", rng.normal(size=(n, r, c))) + image("PIXEL_GRID_FITS", rng.normal(size=(n, r, c)))
            + image("VALIDATION_STARS", rng.normal(size=(3, r, c)))
            + table("FLAGS", [("FLAG_SNR", [1.0, 2.0]), ("FLAG_S_BLACKLIST_MODEL", [4.0]),
                              ("FLAG_PIXEL_GRID_OR_S_BLACKLIST_LEARNED", [])])
            + table("ANALYTIC_PROFILE", [("s_MATRIX", field[0]), ("g1_MATRIX", field[1]), ("g2_MATRIX", field[2])]))
    path = tmp_path / "full_summary.shopt"
    path.write_bytes(blob)
    poly, deg, got_field = stage.read_summary(path)
    assert deg == degree and np.array_equal(poly, cube.transpose(2, 1, 0)) and np.array_equal(got_field, field)


def test_full_layout_writer(stage, tmp_path):
    """shopt_summary_write_full against dataOutprocessing.jl:86-168 read through the independent parser: HDU order and
    names, the ragged zero-padded tables, and the axis order of each cube (VIGNETS is the stacked (r, c, N) array,
    PIXEL_GRID_FITS / VALIDATION_STARS are numpy stacks (N, r, c) of the Julia matrices)."""
    rng = np.random.default_rng(5)
    degree, r, c, n, nv, nt = 2, 5, 4, 9, 3, 6
    K = 6
    poly = rng.normal(size=(K, c, r))
    field = rng.normal(size=(3, 10))
    u, v = rng.normal(size=n), rng.normal(size=n)
    model, learned = rng.normal(size=(3, n)), rng.normal(size=(3, n))
    stars, pgf, vstars = rng.normal(size=(n, c, r)), rng.normal(size=(n, c, r)), rng.normal(size=(nv, c, r))
    tu, tv, vu, vv = rng.normal(size=nt), rng.normal(size=nt), rng.normal(size=nv), rng.normal(size=nv)
    path = tmp_path / "full.shopt"
    stage.write_summary_full(path, poly, degree, field, u, v, model, learned, stars, pgf, tu, tv, vstars, vu, vv,
                             flag_snr=[3.0, 8.0, 11.0], flag_blacklist=[2.0], flag_failed=[])
    hdus = parse_fits(path)
    assert [h["EXTNAME"] for h, _ in hdus] == ["POLYNOMIAL_MATRIX", "SUMMARY_STATISTICS", "VIGNETS", "PIXEL_GRID_FITS",
                                               "VALIDATION_STARS", "FLAGS", "ANALYTIC_PROFILE"]
    assert [h.get("XTENSION", "PRIMARY") for h, _ in hdus] == ["PRIMARY", "BINTABLE", "IMAGE", "IMAGE", "IMAGE", "BINTABLE",
                                                                "BINTABLE"]
    h0, d0 = hdus[0]
    assert np.array_equal(np.frombuffer(d0, dtype=">f8").reshape(r, c, K), poly.transpose(2, 1, 0))
    h1, d1 = hdus[1]
    names = ["U_COORDINATES", "V_COORDINATES", "TRAINING_U_COORDS", "TRAINING_V_COORDS", "VALIDATION_U_COORDS",
             "VALIDATION_V_COORDS", "S_MODEL", "g1_MODEL", "g2_MODEL", "s_DATA", "g1_DATA", "g2_DATA", "POLYNOMIAL_DEGREE"]
    assert [h1["TTYPE%d" % (i + 1)] for i in range(13)] == names and int(h1["NAXIS2"]) == n and int(h1["NAXIS1"]) == 104
    tab = np.frombuffer(d1, dtype=">f8").reshape(n, 13)
    pad = lambda a: np.concatenate([a, np.zeros(n - len(a))])
    want = [u, v, tu, tv, vu, vv, model[0], model[1], model[2], learned[0], learned[1], learned[2], np.array([float(degree)])]
    for j, w in enumerate(want):
        assert np.array_equal(tab[:, j], pad(w)), names[j]
    # numpy shapes as astropy would report them (reversed NAXISn)
    shape = lambda h: tuple(int(h["NAXIS%d" % a]) for a in (3, 2, 1))
    assert shape(hdus[2][0]) == (r, c, n) and shape(hdus[3][0]) == (n, r, c) and shape(hdus[4][0]) == (nv, r, c)
    vig = np.frombuffer(hdus[2][1], dtype=">f8").reshape(r, c, n)
    assert np.array_equal(vig, stars.transpose(2, 1, 0))                   # starCatalog[a, b, i]
    pg = np.frombuffer(hdus[3][1], dtype=">f8").reshape(n, r, c)
    assert np.array_equal(pg, pgf.transpose(0, 2, 1))                      # pixelGridFits[i][a, b]
    va = np.frombuffer(hdus[4][1], dtype=">f8").reshape(nv, r, c)
    assert np.array_equal(va, vstars.transpose(0, 2, 1))
    h5, d5 = hdus[5]
    assert [h5["TTYPE%d" % i] for i in (1, 2, 3)] == ["FLAG_SNR", "FLAG_S_BLACKLIST_MODEL",
                                                      "FLAG_PIXEL_GRID_OR_S_BLACKLIST_LEARNED"]
    fl = np.frombuffer(d5, dtype=">f8").reshape(3, 3)
    assert np.array_equal(fl, np.array([[3.0, 2.0, 0.0], [8.0, 0.0, 0.0], [11.0, 0.0, 0.0]]))
    h6, d6 = hdus[6]
    assert np.array_equal(np.frombuffer(d6, dtype=">f8").reshape(10, 3).T, field)
    # and the library's reader takes the file back (read_shopt: f[0], f[1]['POLYNOMIAL_DEGREE'][0], f[6])
    got_poly, got_deg, got_field = stage.read_summary(path)
    assert got_deg == degree and np.array_equal(got_poly, poly) and np.array_equal(got_field, field)
    # an empty catalogue still writes a well-formed file
    stage.write_summary_full(tmp_path / "empty.shopt", poly, degree, field, [], [], np.zeros((3, 0)), np.zeros((3, 0)),
                             np.zeros((0, c, r)), np.zeros((0, c, r)), [], [], np.zeros((0, c, r)), [], [])
    assert len(parse_fits(tmp_path / "empty.shopt")) == 7


def test_bad_files_are_refused(stage, tmp_path):
    import pytest
    from shopt_b200 import _lib
    p = tmp_path / "junk.shopt"
    p.write_bytes(b"not a fits file")
    with pytest.raises(_lib.ShoptError):
        stage.read_summary(p)
    with pytest.raises(_lib.ShoptError):
        stage.read_summary(tmp_path / "missing.shopt")
