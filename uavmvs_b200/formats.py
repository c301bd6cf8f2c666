"""On-disk formats either side of the hot path, restated from the
reference's readers / writers so that its files round-trip without MVE:

  trajectory text   libs/utp/trajectory_io.cpp:13-73   N, then per view: pos(3) rot(9, row-major) flen
  PLY mesh / cloud  libs/util/io.h:49-63, libs/util/cio.h:15-50 (MVE's PLY: x y z [nx ny nz] [value]
                    [confidence], faces as `list uchar int vertex_indices`)
  .npy output       libs/util/numpy_io.h:59-102 (format 2.0 header, 16-byte aligned, 2-D shape)
  VOL container     libs/geom/volume_io.h:22-107 (guidance volume: sparse list of float images)

Host-side helpers only (numpy); nothing here touches the device.  MVE itself is not in the reference
tree, so byte-level parity of the TEXT writers' whitespace is unpinned; every reader accepts what the
reference's readers accept (whitespace-separated tokens).
"""
import io
import struct

import numpy as np

# ------------------------------------------------------------------------------------ trajectory


def load_trajectory(path):
    """-> (pos [V,3], rot [V,3,3] world-to-camera rotation, flen [V]) as float32
    (libs/utp/trajectory_io.cpp:37-73: the reader consumes whitespace-separated numbers)."""
    tok = open(path).read().split()
    if not tok:
        raise ValueError("Invalid trajectory file")
    n = int(tok[0])
    if len(tok) < 1 + 13 * n:
        raise ValueError("Invalid trajectory file")
    a = np.array(tok[1:1 + 13 * n], dtype=np.float32).reshape(n, 13)
    return a[:, :3].copy(), a[:, 3:12].reshape(n, 3, 3).copy(), a[:, 12].copy()


def save_trajectory(path, pos, rot, flen):
    """libs/utp/trajectory_io.cpp:13-35: length, then `pos`, the 3x3 `rot` (one row per line), `flen`."""
    pos, rot, flen = np.asarray(pos, np.float32), np.asarray(rot, np.float32).reshape(-1, 3, 3), np.asarray(flen, np.float32)
    with open(path, "w") as f:
        f.write("%d\n" % pos.shape[0])
        for p, r, fl in zip(pos, rot, flen):
            f.write(" ".join(repr(float(x)) for x in p) + "\n")
            for row in r:
                f.write(" ".join(repr(float(x)) for x in row) + "\n")
            f.write(repr(float(fl)) + "\n")


def views_from_trajectory(pos, rot, flen, width=1920, height=1080):
    """The kernel arguments the apps derive from every mve::CameraInfo
    (apps/evaluate_trajectory/evaluate_trajectory.cu:155-157): view_pos, w2c = [R | -R pos],
    calib = fill_calibration(width, height) with principal point (0.5, 0.5), pixel aspect 1
    -> float32 [V, 28] rows (pos, w2c row-major, calib row-major) as the C ABI takes them."""
    pos = np.asarray(pos, np.float32)
    rot = np.asarray(rot, np.float32).reshape(-1, 3, 3)
    flen = np.asarray(flen, np.float32)
    V = pos.shape[0]
    out = np.zeros((V, 28), np.float32)
    out[:, :3] = pos
    w2c = np.zeros((V, 4, 4), np.float32)
    w2c[:, :3, :3] = rot
    w2c[:, :3, 3] = -np.einsum("vij,vj->vi", rot, pos)            # trans = -rot * pos (:60)
    w2c[:, 3, 3] = 1.0
    out[:, 3:19] = w2c.reshape(V, 16)
    ax = flen * np.float32(width if width >= height else height)   # landscape / portrait branch of fill_calibration
    K = np.zeros((V, 3, 3), np.float32)
    K[:, 0, 0], K[:, 1, 1] = ax, ax
    K[:, 0, 2], K[:, 1, 2], K[:, 2, 2] = width * 0.5, height * 0.5, 1.0
    out[:, 19:28] = K.reshape(V, 9)
    return out


# ------------------------------------------------------------------------------------ PLY
_PLY_TYPES = {"char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1", "short": "i2", "int16": "i2",
              "ushort": "u2", "uint16": "u2", "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4",
              "float": "f4", "float32": "f4", "double": "f8", "float64": "f8"}


def _ply_header(f):
    if f.readline().strip() != b"ply":
        raise ValueError("not a PLY file")
    fmt, elements = None, []
    while True:
        line = f.readline()
        if not line:
            raise ValueError("PLY header not terminated")
        t = line.decode("ascii", "replace").split()
        if not t or t[0] == "comment" or t[0] == "obj_info":
            continue
        if t[0] == "format":
            fmt = t[1]
        elif t[0] == "element":
            elements.append({"name": t[1], "count": int(t[2]), "props": []})
        elif t[0] == "property":
            if t[1] == "list":
                elements[-1]["props"].append(("list", _PLY_TYPES[t[2]], _PLY_TYPES[t[3]], t[4]))
            else:
                elements[-1]["props"].append(("scalar", _PLY_TYPES[t[1]], None, t[2]))
        elif t[0] == "end_header":
            break
    if fmt not in ("ascii", "binary_little_endian", "binary_big_endian"):
        raise ValueError("unsupported PLY format %r" % fmt)
    return fmt, elements


def _read_element(f, fmt, el):
    """-> dict name -> array (scalars) or list of arrays (list properties)."""
    n, props = el["count"], el["props"]
    end = "<" if fmt != "binary_big_endian" else ">"
    if all(p[0] == "scalar" for p in props):
        if fmt == "ascii":
            rows = [f.readline().split() for _ in range(n)]
            cols = {p[3]: np.array([r[i] for r in rows], dtype=np.dtype(p[1])) for i, p in enumerate(props)}
            return cols
        dt = np.dtype([(p[3], end + p[1]) for p in props])
        a = np.frombuffer(f.read(dt.itemsize * n), dt, n)
        return {p[3]: a[p[3]].astype(np.dtype(p[1])) for p in props}
    out = {p[3]: [] for p in props}
    for _ in range(n):
        if fmt == "ascii":
            t = f.readline().split()
            k = 0
            for p in props:
                if p[0] == "list":
                    c = int(t[k]); out[p[3]].append(np.array(t[k + 1:k + 1 + c], dtype=np.dtype(p[2]))); k += 1 + c
                else:
                    out[p[3]].append(np.dtype(p[1]).type(t[k])); k += 1
        else:
            for p in props:
                if p[0] == "list":
                    c = int(np.frombuffer(f.read(np.dtype(p[1]).itemsize), end + p[1])[0])
                    out[p[3]].append(np.frombuffer(f.read(np.dtype(p[2]).itemsize * c), end + p[2]).astype(np.dtype(p[2])))
                else:
                    out[p[3]].append(np.frombuffer(f.read(np.dtype(p[1]).itemsize), end + p[1])[0])
    return out


def load_ply(path):
    """Vertices, optional normals / value / confidence, triangle faces (MVE's names):
    -> dict(vertices [n,3] f32, normals or None, values or None, confidences or None, faces [m,3] u32)."""
    with open(path, "rb") as f:
        fmt, elements = _ply_header(f)
        data = {}
        for el in elements:
            # fast path: the usual `list uchar int vertex_indices` triangle element in binary files
            if (fmt == "binary_little_endian" and len(el["props"]) == 1 and el["props"][0][0] == "list"):
                p = el["props"][0]
                pos = f.tell()
                dt = np.dtype([("n", "<" + p[1]), ("i", "<" + p[2], 3)])
                raw = f.read(dt.itemsize * el["count"])
                a = np.frombuffer(raw, dt, el["count"]) if len(raw) == dt.itemsize * el["count"] else None
                if a is not None and (el["count"] == 0 or np.all(a["n"] == 3)):
                    data[el["name"]] = {p[3]: a["i"].astype(np.dtype(p[2]))}
                    continue
                f.seek(pos)
            data[el["name"]] = _read_element(f, fmt, el)
    v = data.get("vertex", {})
    if not all(k in v for k in ("x", "y", "z")):
        raise ValueError("PLY file has no vertex positions")
    out = {"vertices": np.stack([v["x"], v["y"], v["z"]], 1).astype(np.float32)}
    out["normals"] = np.stack([v["nx"], v["ny"], v["nz"]], 1).astype(np.float32) if "nx" in v else None
    out["values"] = np.asarray(v["value"], np.float32) if "value" in v else None
    out["confidences"] = np.asarray(v["confidence"], np.float32) if "confidence" in v else None
    faces = np.zeros((0, 3), np.uint32)
    fe = data.get("face")
    if fe:
        key = "vertex_indices" if "vertex_indices" in fe else ("vertex_index" if "vertex_index" in fe else None)
        if key is not None:
            fl = fe[key]
            if isinstance(fl, np.ndarray):
                faces = fl.reshape(-1, 3).astype(np.uint32)
            else:
                if any(len(x) != 3 for x in fl):
                    raise ValueError("only triangle meshes are supported")
                faces = np.array(fl, np.uint32).reshape(-1, 3)
    out["faces"] = faces
    return out


def save_ply(path, vertices, faces=None, normals=None, values=None, confidences=None):
    """Binary little-endian PLY with MVE's property names (what load_point_cloud / load_ply_mesh read)."""
    vertices = np.asarray(vertices, np.float32).reshape(-1, 3)
    cols, names = [vertices], ["x", "y", "z"]
    if normals is not None:
        cols.append(np.asarray(normals, np.float32).reshape(-1, 3)); names += ["nx", "ny", "nz"]
    if values is not None:
        cols.append(np.asarray(values, np.float32).reshape(-1, 1)); names += ["value"]
    if confidences is not None:
        cols.append(np.asarray(confidences, np.float32).reshape(-1, 1)); names += ["confidence"]
    tab = np.ascontiguousarray(np.concatenate(cols, 1).astype("<f4"))
    faces = np.zeros((0, 3), np.uint32) if faces is None else np.asarray(faces).reshape(-1, 3)
    with open(path, "wb") as f:
        f.write(b"ply\nformat binary_little_endian 1.0\n")
        f.write(b"element vertex %d\n" % vertices.shape[0])
        for nme in names:
            f.write(b"property float %s\n" % nme.encode())
        f.write(b"element face %d\nproperty list uchar int vertex_indices\nend_header\n" % faces.shape[0])
        f.write(tab.tobytes())
        rec = np.empty(faces.shape[0], np.dtype([("n", "u1"), ("i", "<i4", 3)]))
        rec["n"] = 3
        rec["i"] = faces.astype(np.int32)
        f.write(rec.tobytes())


def load_point_cloud(path):
    """libs/util/cio.h:15-50: vertices and normals are required; values default to 0, qualities
    (the PLY `confidence`) to 1.  -> (points, normals, values, qualities)."""
    d = load_ply(path)
    if d["normals"] is None:
        raise ValueError("Point cloud has no vertex normals.")
    n = d["vertices"].shape[0]
    values = d["values"] if d["values"] is not None else np.zeros(n, np.float32)
    quals = d["confidences"] if d["confidences"] is not None else np.ones(n, np.float32)
    return d["vertices"], d["normals"], values, quals


# ------------------------------------------------------------------------------------ .npy (numpy_io.h)
def save_numpy_file(path, *vectors):
    """libs/util/numpy_io.h:59-102: m equally long vectors of one arithmetic type become the ROWS of an
    (m, n) C-order array in a format-2.0 .npy file whose header is padded with spaces to a multiple of 16."""
    vs = [np.ascontiguousarray(v).reshape(-1) for v in vectors]
    if not vs or any(v.shape != vs[0].shape or v.dtype != vs[0].dtype for v in vs):
        raise ValueError("vectors must share length and type")
    dt = vs[0].dtype
    kind = "f" if dt.kind == "f" else "i"
    d = "{'descr': '<%s%d', 'fortran_order': False, 'shape': (%d, %d), }" % (kind, dt.itemsize, len(vs), vs[0].shape[0])
    sig = b"\x93NUMPY"
    header_length = len(d) + len(sig) + 6 + 1
    padding = 16 * ((header_length + 15) // 16) - header_length
    header_length += padding - len(sig) - 6
    with open(path, "wb") as f:
        f.write(sig + b"\x02\x00" + struct.pack("<I", header_length) + d.encode() + b" " * padding + b"\n")
        for v in vs:
            f.write(v.astype(dt.newbyteorder("<")).tobytes())


# ------------------------------------------------------------------------------------ VOL (volume_io.h)
def save_volume(path, width, height, depth, minimum, maximum, images):
    """libs/geom/volume_io.h:22-46.  images: dict linear index -> float32 array [h, w] or [h, w, c]."""
    with open(path, "wb") as f:
        f.write(b"VOL 0.1\n%d %d %d \n" % (width, height, depth))
        # `out << float` in the C++ default format (precision 6, %g); MVE prints a vector as "x y z" (volume_io.h:34)
        f.write((" ".join("%g" % float(np.float32(x)) for x in minimum) + " " + " ".join("%g" % float(np.float32(x)) for x in maximum)).encode())
        for i in sorted(images):
            img = np.asarray(images[i], np.float32)
            if img.ndim == 2:
                img = img[:, :, None]
            f.write(b"\n%d %d %d %d\n" % (i, img.shape[1], img.shape[0], img.shape[2]))
            f.write(np.ascontiguousarray(img).astype("<f4").tobytes())


def load_volume(path):
    """libs/geom/volume_io.h:48-105 -> (width, height, depth, minimum[3], maximum[3], {index: image [h, w, c]})."""
    raw = open(path, "rb").read()
    f = io.BytesIO(raw)
    head = f.readline().split()
    if not head or head[0] != b"VOL":
        raise ValueError("Not a Volume file")
    if len(head) < 2 or head[1] != b"0.1":
        raise ValueError("Incompatible version of Volume file")
    dims = f.readline().split()
    width, height, depth = (int(x) for x in dims[:3])
    # the bounds line has no terminator of its own: it ends where the first image record (or the file) begins
    pos = f.tell()
    nl = raw.find(b"\n", pos)
    line = raw[pos:] if nl < 0 else raw[pos:nl]
    b = [float(x) for x in line.split()]
    if len(b) != 6:
        raise ValueError("Corrupt Volume file header")
    f.seek(len(raw) if nl < 0 else nl + 1)
    images = {}
    while True:
        rec = f.readline()
        if not rec.strip():
            if not rec:
                break
            continue
        idx, iw, ih, ic = (int(x) for x in rec.split()[:4])
        nbytes = 4 * iw * ih * ic
        buf = f.read(nbytes)
        if len(buf) != nbytes:
            raise ValueError("Corrupt Volume file")
        images[idx] = np.frombuffer(buf, "<f4").reshape(ih, iw, ic).copy()
        f.read(1) if raw[f.tell():f.tell() + 1] == b"\n" else None
    return width, height, depth, np.array(b[:3], np.float32), np.array(b[3:], np.float32), images
