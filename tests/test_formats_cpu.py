"""f3: PCD reader (ascii / binary / binary_compressed) and GraspListMsg serialisation — host code of
libag2.so, exercised without a GPU against files written by an independent Python writer."""
import struct

import numpy as np

from agile_grasp2_b200 import capi


def _header(n, data, with_normals, rgb_float=False):
    f, s, t, c = ["x", "y", "z"], ["4"] * 3, ["F"] * 3, ["1"] * 3
    f.append("rgb" if rgb_float else "rgba"); s.append("4"); t.append("F" if rgb_float else "U"); c.append("1")
    if with_normals:
        f += ["normal_x", "normal_y", "normal_z", "curvature"]; s += ["4"] * 4; t += ["F"] * 4; c += ["1"] * 4
    return ("# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS %s\nSIZE %s\nTYPE %s\nCOUNT %s\n"
            "WIDTH %d\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS %d\nDATA %s\n"
            % (" ".join(f), " ".join(s), " ".join(t), " ".join(c), n, n, data)).encode()


def _lzf_literal(raw: bytes) -> bytes:
    """A valid LZF stream made of literal runs plus one back-reference (exercises both opcodes)."""
    out = bytearray()
    head = raw[:40]
    body = raw[40:]
    for i in range(0, len(head), 32):
        chunk = head[i:i + 32]
        out.append(len(chunk) - 1); out += chunk
    # back reference if the next 3 bytes repeat something 8 bytes back; otherwise literals only
    i = 0
    pos = len(head)
    while i < len(body):
        if pos >= 8 and body[i:i + 3] == raw[pos - 8:pos - 5] and len(body) - i >= 3:
            out.append((1 << 5) | 0); out.append(7)      # len 3, offset 8
            i += 3; pos += 3
            continue
        chunk = body[i:i + 32]
        # do not let a literal run cross a point where a back reference is possible: keep it simple
        out.append(len(chunk) - 1); out += chunk
        i += len(chunk); pos += len(chunk)
    return bytes(out)


def _cloud(n=257, seed=0):
    rng = np.random.default_rng(seed)
    xyz = rng.normal(0, 1, (n, 3)).astype(np.float32)
    xyz[3] = np.nan
    nrm = rng.normal(0, 1, (n, 3)).astype(np.float32)
    rgba = rng.integers(0, 2 ** 32, n, dtype=np.uint64).astype(np.uint32)
    return xyz, nrm, rgba


def test_pcd_binary_and_compressed(tmp_path):
    xyz, nrm, rgba = _cloud()
    n = len(xyz)
    curv = np.zeros(n, np.float32)
    rec = np.zeros(n, dtype=[("xyz", "<f4", 3), ("rgba", "<u4"), ("n", "<f4", 3), ("c", "<f4")])
    rec["xyz"], rec["rgba"], rec["n"], rec["c"] = xyz, rgba, nrm, curv
    p = tmp_path / "b.pcd"
    p.write_bytes(_header(n, "binary", True) + rec.tobytes())
    gx, gn, gc = capi.pcd_read(str(p))
    assert np.array_equal(gx.view(np.uint32), xyz.view(np.uint32))
    assert np.array_equal(gn, nrm) and np.array_equal(gc, rgba)
    # binary_compressed: column-wise fields, LZF
    soa = b"".join([xyz[:, 0].tobytes(), xyz[:, 1].tobytes(), xyz[:, 2].tobytes(), rgba.tobytes(),
                    nrm[:, 0].tobytes(), nrm[:, 1].tobytes(), nrm[:, 2].tobytes(), curv.tobytes()])
    comp = _lzf_literal(soa)
    p = tmp_path / "c.pcd"
    p.write_bytes(_header(n, "binary_compressed", True) + struct.pack("<II", len(comp), len(soa)) + comp)
    gx, gn, gc = capi.pcd_read(str(p))
    assert np.array_equal(gx.view(np.uint32), xyz.view(np.uint32))
    assert np.array_equal(gn, nrm) and np.array_equal(gc, rgba)


def test_pcd_ascii_and_own_writer(tmp_path):
    xyz, nrm, rgba = _cloud(100, 1)
    lines = [" ".join(["%.9g" % v for v in xyz[i]] + [str(int(rgba[i]))]) for i in range(len(xyz))]
    p = tmp_path / "a.pcd"
    p.write_bytes(_header(len(xyz), "ascii", False) + ("\n".join(lines) + "\n").encode())
    gx, gn, gc = capi.pcd_read(str(p))
    assert gn is None
    assert np.array_equal(gx.view(np.uint32), xyz.view(np.uint32))  # %.9g round-trips float32
    assert np.array_equal(gc, rgba)
    for binary in (True, False):
        q = tmp_path / ("w%d.pcd" % binary)
        capi.pcd_write(str(q), xyz, nrm, rgba, binary=binary)
        gx, gn, gc = capi.pcd_read(str(q))
        assert np.array_equal(gx.view(np.uint32), xyz.view(np.uint32))
        assert np.array_equal(gn, nrm) and np.array_equal(gc, rgba)


def test_pcd_errors(tmp_path):
    import pytest
    with pytest.raises(capi.Ag2Error):
        capi.pcd_read(str(tmp_path / "missing.pcd"))
    (tmp_path / "bad.pcd").write_bytes(b"FIELDS a b\nSIZE 4 4\nTYPE F F\nCOUNT 1 1\nPOINTS 1\nDATA ascii\n1 2\n")
    with pytest.raises(capi.Ag2Error):
        capi.pcd_read(str(tmp_path / "bad.pcd"))


def test_grasp_list_wire_format():
    h = np.zeros(2, capi.HYPOTHESIS_DTYPE)
    rng = np.random.default_rng(3)
    for f in ("surface", "bottom", "top", "axis", "approach", "binormal"):
        h[f] = rng.normal(0, 1, (2, 3))
    h["width"], h["score"] = [0.05, 0.061], [512.25, -3.5]
    b = capi.serialize_grasp_list(h, seq=7, sec=11, nsec=13, frame_id="base")
    assert len(b) == 12 + 4 + 4 + 4 + 2 * 152
    seq, sec, nsec, fl = struct.unpack_from("<IIII", b, 0)
    assert (seq, sec, nsec, fl) == (7, 11, 13, 4) and b[16:20] == b"base"
    (cnt,) = struct.unpack_from("<I", b, 20)
    assert cnt == 2
    off = 24
    for i in range(2):
        v = struct.unpack_from("<18d2f", b, off)
        exp = np.concatenate([h[f][i] for f in ("surface", "bottom", "top", "axis", "approach", "binormal")])
        assert np.array_equal(np.array(v[:18]), exp)
        assert v[18] == np.float32(h["width"][i]) and v[19] == np.float32(h["score"][i])
        off += 152


def test_pcd_header_variants(tmp_path):
    """Headers real .pcd files carry and pcl::io::loadPCDFile accepts (cloud_camera.cpp:231-240): double-typed
    coordinates, organised clouds (WIDTH x HEIGHT), fields with COUNT > 1 and extra fields around x y z, a v0.6 header
    without COUNT, no VIEWPOINT line, CRLF line ends, PCL's "_" padding field, ascii data."""
    rng = np.random.default_rng(1)
    n = 12
    xyz = rng.normal(0, 1, (n, 3))
    f32 = xyz.astype(np.float32)
    tail = "WIDTH %d\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS %d\nDATA binary\n" % (n, n)
    cases = {
        "double": ("VERSION 0.7\nFIELDS x y z\nSIZE 8 8 8\nTYPE F F F\nCOUNT 1 1 1\n" + tail, xyz.astype("<f8").tobytes()),
        "organised": ("VERSION 0.7\nFIELDS x y z\nSIZE 4 4 4\nTYPE F F F\nCOUNT 1 1 1\nWIDTH 4\nHEIGHT 3\n"
                      "VIEWPOINT 0 0 0 1 0 0 0\nPOINTS 12\nDATA binary\n", f32.tobytes()),
        "count5": ("VERSION 0.7\nFIELDS fpfh x y z intensity\nSIZE 4 4 4 4 2\nTYPE F F F F U\nCOUNT 5 1 1 1 1\n" + tail,
                   b"".join(struct.pack("<5f3fH", 0, 1, 2, 3, 4, *f32[i], 7) for i in range(n))),
        "ascii_no_viewpoint": ("# comment\nVERSION .7\nFIELDS x y z rgb\nSIZE 4 4 4 4\nTYPE F F F F\nCOUNT 1 1 1 1\n"
                               "WIDTH 12\nHEIGHT 1\nPOINTS 12\nDATA ascii\n",
                               ("\n".join("%r %r %r 4.2e-38" % tuple(float(v) for v in f32[i]) for i in range(n)) + "\n").encode()),
        "v06_no_count": ("VERSION 0.6\nFIELDS x y z\nSIZE 4 4 4\nTYPE F F F\nWIDTH 12\nHEIGHT 1\nPOINTS 12\nDATA binary\n",
                         f32.tobytes()),
        "crlf": (("VERSION 0.7\nFIELDS x y z\nSIZE 4 4 4\nTYPE F F F\nCOUNT 1 1 1\n" + tail).replace("\n", "\r\n"), f32.tobytes()),
        "padding": ("VERSION 0.7\nFIELDS x y z _ rgba\nSIZE 4 4 4 1 4\nTYPE F F F U U\nCOUNT 1 1 1 4 1\n" + tail,
                    b"".join(struct.pack("<3f4BI", *f32[i], 0, 0, 0, 0, 0xFF112233) for i in range(n))),
    }
    for name, (hdr, body) in cases.items():
        path = tmp_path / (name + ".pcd")
        path.write_bytes(hdr.encode() + body)
        got = capi.pcd_read(str(path))
        got_xyz = got[0] if isinstance(got, tuple) else got["xyz"]
        assert got_xyz.shape == (n, 3), name
        assert np.array_equal(got_xyz, f32), name


# ---- .caffemodel (protobuf wire format of caffe.NetParameter), written here by hand -------------------------------

def _varint(v: int) -> bytes:
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        out.append(b | (0x80 if v else 0))
        if not v:
            return bytes(out)


def _field(num: int, wire: int, payload: bytes) -> bytes:
    return _varint((num << 3) | wire) + (_varint(len(payload)) + payload if wire == 2 else payload)


def _blob(a: np.ndarray, packed: bool, legacy_dims: bool) -> bytes:
    a = np.ascontiguousarray(a, "<f4")
    out = b""
    if legacy_dims:  # BlobProto.num / channels / height / width = 1..4
        dims = (list(a.shape) + [1, 1, 1, 1])[:4]
        out += b"".join(_field(i + 1, 0, _varint(d)) for i, d in enumerate(dims))
    else:  # BlobProto.shape = 7 { dim = 1, packed }
        out += _field(7, 2, _field(1, 2, b"".join(_varint(d) for d in a.shape)))
    if packed:
        out += _field(5, 2, a.tobytes())
    else:
        out += b"".join(_field(5, 5, a.reshape(-1)[i:i + 1].tobytes()) for i in range(a.size))
    return out


def _caffemodel(w: dict, v1: bool, packed: bool = True, skip_bias_of=(), order=("conv1", "conv2", "ip1", "ip2")) -> bytes:
    """NetParameter: name = 1; V2 layer = 100 {name = 1, type = 2 (string), bottom = 3, top = 4, blobs = 7};
    V1 layers = 2 {bottom = 2, top = 3, name = 4, type = 5 (enum), blobs = 6}; plus layers without blobs."""
    name_f, blobs_f, layer_f = (4, 6, 2) if v1 else (1, 7, 100)
    out = _field(1, 2, b"LeNet")
    pool = _field(name_f, 2, b"pool1") + (_field(5, 0, _varint(17)) if v1 else _field(2, 2, b"Pooling"))
    out += _field(layer_f, 2, pool)
    for name in order:
        body = _field(name_f, 2, name.encode())
        body += _field(5, 0, _varint(4)) if v1 else _field(2, 2, b"Convolution")
        body += _field(2 if v1 else 3, 2, b"data") + _field(3 if v1 else 4, 2, name.encode())
        # conv / ip weights are unpacked only for the small layers (an unpacked ip1 would be 18 MB of 5-byte fields)
        body += _field(blobs_f, 2, _blob(w[name + "_w"], packed or name == "ip1", v1))
        if name not in skip_bias_of:
            body += _field(blobs_f, 2, _blob(w[name + "_b"], packed, v1))
        out += _field(layer_f, 2, body)
    return out


def _weights(seed=3):
    rng = np.random.default_rng(seed)
    return {k: rng.normal(0, 0.1, shp).astype(np.float32) for k, shp in capi.LENET_SHAPES}


def test_caffemodel_v2_v1_packed_and_repeated(tmp_path):
    """Net::CopyTrainedLayersFrom's input (caffe_classifier.cpp:13-14): current `layer` and legacy V1 `layers` messages,
    packed and one-float-per-field blob data, blobs with shape or with num/channels/height/width, layers in any order,
    layers without blobs in between -- all give back the blobs bit for bit."""
    w = _weights()
    for v1, packed, order in ((False, True, ("conv1", "conv2", "ip1", "ip2")), (True, True, ("ip2", "conv2", "ip1", "conv1")),
                              (False, False, ("conv1", "conv2", "ip1", "ip2")), (True, False, ("conv1", "conv2", "ip1", "ip2"))):
        path = tmp_path / ("lenet_%d_%d.caffemodel" % (v1, packed))
        path.write_bytes(_caffemodel(w, v1, packed, order=order))
        got = capi.read_lenet_caffemodel(str(path))
        for k, _ in capi.LENET_SHAPES:
            assert np.array_equal(got[k], w[k]), (v1, packed, k)


def test_caffemodel_missing_bias_and_errors(tmp_path):
    import pytest

    w = _weights(4)
    path = tmp_path / "nobias.caffemodel"
    path.write_bytes(_caffemodel(w, False, skip_bias_of=("conv2",)))
    got = capi.read_lenet_caffemodel(str(path))
    assert np.array_equal(got["conv2_w"], w["conv2_w"]) and not got["conv2_b"].any()  # bias_term: false -> zeros
    good = _caffemodel(w, False)
    bad = dict(w)
    bad["conv1_w"] = np.zeros((20, 1, 5, 5), np.float32)  # a grey-scale LeNet: not the prototxt's shapes
    cases = {
        "truncated": good[: len(good) // 2],
        "wrong_shape": _caffemodel(bad, False),
        "missing_layer": _caffemodel(w, False, order=("conv1", "conv2", "ip1")),
        "garbage": bytes(range(256)) * 7,
    }
    for name, data in cases.items():
        p = tmp_path / (name + ".caffemodel")
        p.write_bytes(data)
        with pytest.raises(capi.Ag2Error):
            capi.read_lenet_caffemodel(str(p))
    with pytest.raises(capi.Ag2Error):
        capi.read_lenet_caffemodel(str(tmp_path / "does_not_exist.caffemodel"))
