"""Test infrastructure: a minimal glTF 2.0 / GLB writer (pure Python) used to feed
`gpu_raytracer_b200.load_gltf`, plus an independent numpy restatement of the arithmetic the
reference loader applies (src/gltf_loader.rs) for the expected values."""
import base64
import json
import struct

import numpy as np

F32 = np.float32


class BinBuilder:
    def __init__(self):
        self.blob = bytearray()
        self.views, self.accessors = [], []

    def add(self, array, ctype, atype, stride=None):
        """Append `array` as one bufferView + accessor; returns the accessor index."""
        a = np.ascontiguousarray(array)
        while len(self.blob) % 4:
            self.blob.append(0)
        off = len(self.blob)
        if stride is None:
            self.blob += a.tobytes()
            view = {"buffer": 0, "byteOffset": off, "byteLength": a.nbytes}
        else:  # interleave with padding up to `stride` bytes per element
            esize = a.dtype.itemsize * (a.shape[1] if a.ndim > 1 else 1)
            for row in a.reshape(len(a), -1):
                self.blob += row.tobytes() + b"\xAB" * (stride - esize)
            view = {"buffer": 0, "byteOffset": off, "byteLength": stride * len(a), "byteStride": stride}
        self.views.append(view)
        self.accessors.append({"bufferView": len(self.views) - 1, "componentType": ctype, "count": len(a), "type": atype})
        return len(self.accessors) - 1


def finish(doc, bb, how, path):
    """how: 'glb' | 'bin' (external file) | 'base64'.  Writes the file(s) and returns the main path."""
    doc = dict(doc)
    doc.setdefault("asset", {"version": "2.0"})
    doc["bufferViews"], doc["accessors"] = bb.views, bb.accessors
    blob = bytes(bb.blob)
    if how == "glb":
        doc["buffers"] = [{"byteLength": len(blob)}]
        js = json.dumps(doc).encode()
        js += b" " * (-len(js) % 4)
        bn = blob + b"\0" * (-len(blob) % 4)
        total = 12 + 8 + len(js) + 8 + len(bn)
        out = struct.pack("<4sII", b"glTF", 2, total) + struct.pack("<II", len(js), 0x4E4F534A) + js + \
            struct.pack("<II", len(bn), 0x004E4942) + bn
        p = str(path) + ".glb"
        open(p, "wb").write(out)
        return p
    if how == "bin":
        binp = str(path) + ".bin"
        open(binp, "wb").write(blob)
        doc["buffers"] = [{"byteLength": len(blob), "uri": binp.split("/")[-1]}]
    else:
        doc["buffers"] = [{"byteLength": len(blob), "uri": "data:application/octet-stream;base64," + base64.b64encode(blob).decode()}]
    p = str(path) + ".gltf"
    open(p, "w").write(json.dumps(doc, indent=1))
    return p


# ---- independent restatement of the loader's fp32 arithmetic (glam scalar-math association) ----
def mat_from_trs(t, q, s):
    x, y, z, w = (F32(v) for v in q)
    x2, y2, z2 = x + x, y + y, z + z
    xx2, xy2, xz2, yy2, yz2, zz2 = x2 * x, x2 * y, x2 * z, y2 * y, y2 * z, z2 * z
    sy2, sz2, sx2 = y2 * w, z2 * w, x2 * w
    one = F32(1)
    R = np.identity(4, dtype=F32)  # R[row, col]
    R[0, 0], R[1, 0], R[2, 0] = one - yy2 - zz2, xy2 + sz2, xz2 - sy2
    R[0, 1], R[1, 1], R[2, 1] = xy2 - sz2, one - xx2 - zz2, yz2 + sx2
    R[0, 2], R[1, 2], R[2, 2] = xz2 + sy2, yz2 - sx2, one - xx2 - yy2
    T = np.identity(4, dtype=F32)
    T[:3, 3] = np.asarray(t, F32)
    S = np.diag(np.asarray(list(s) + [1], F32)).astype(F32)
    return mat_mul(mat_mul(T, R), S)


def mat_mul(a, b):
    """glam Mat4 * Mat4: column j = ((a.col0*b0j + a.col1*b1j) + a.col2*b2j) + a.col3*b3j, all fp32."""
    out = np.zeros((4, 4), dtype=F32)
    for j in range(4):
        res = a[:, 0] * b[0, j]
        res = res + a[:, 1] * b[1, j]
        res = res + a[:, 2] * b[2, j]
        res = res + a[:, 3] * b[3, j]
        out[:, j] = res
    return out


def transform_point(m, p):
    p = np.asarray(p, F32)
    res = m[:3, 0] * p[0]
    res = m[:3, 1] * p[1] + res
    res = m[:3, 2] * p[2] + res
    return (m[:3, 3] + res).astype(F32)


def transform_vector(m, p):
    p = np.asarray(p, F32)
    res = m[:3, 0] * p[0]
    res = m[:3, 1] * p[1] + res
    res = m[:3, 2] * p[2] + res
    return res.astype(F32)


def normalize(v):
    v = np.asarray(v, F32)
    ln = np.sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2], dtype=F32)
    return (v * (F32(1) / ln)).astype(F32)


def dedup_triangles(positions, index_triples, start_vertex=0):
    """Per-primitive bit-exact position de-duplication in first-use order (gltf_loader.rs:308-334)."""
    seen, verts, tris = {}, [], []
    for tri in index_triples:
        out = []
        for i in tri:
            key = positions[i].astype(F32).tobytes()
            if key not in seen:
                seen[key] = start_vertex + len(verts)
                verts.append(positions[i].astype(F32))
            out.append(seen[key])
        tris.append(out)
    return np.array(verts, dtype=F32).reshape(-1, 3), np.array(tris, dtype=np.uint32).reshape(-1, 3)
