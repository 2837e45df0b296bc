"""The reference's on-disk model format: the `point_cloud.ply` written by GaussianModel.save_ply and read by
GaussianModel.load_ply (scene/gaussian_model.py:277-364), without the `plyfile` package (absent from this image).

File layout (what plyfile's `PlyData([PlyElement.describe(elements, 'vertex')]).write(path)` emits for the reference's
structured array, scene/gaussian_model.py:290-296):

    ply
    format binary_little_endian 1.0
    element vertex <P>
    property float x ... property float rot_3        one line per attribute, all `float` (numpy 'f4')
    end_header
    <P records of 4-byte little-endian floats, attributes interleaved per vertex>

Attribute order (construct_list_of_attributes, :263-275): x y z | nx ny nz (zeros) | f_dc_0..2 | f_rest_0..3(K-1)-1 |
opacity | scale_0..2 | rot_0..3.  The SH blocks are stored CHANNEL-major: the model keeps features as (P, coeffs, 3) and
save_ply writes `transpose(1, 2).flatten(1)` (:281-282), so f_rest_j = rest[:, j % (K-1), j // (K-1)]; load_ply undoes it
(:334-336, :352-353).  Values are the raw (pre-activation) parameters: log-scales, logit opacity, unnormalised quaternion.

`save_ply(path, ...)` / `load_ply(path, max_sh_degree)` work on numpy arrays or torch tensors (any device) and return
numpy float32; `gaussians_from_ply` returns torch tensors on a device, shaped as the model's parameters.  The reader
also accepts ASCII and big-endian files and other scalar property types, as plyfile does; list properties are rejected.
"""
import os

import numpy as np

_PLY_TYPES = {  # PLY scalar type names (both spellings) -> numpy type codes
    "char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1", "short": "i2", "int16": "i2", "ushort": "u2", "uint16": "u2",
    "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4", "float": "f4", "float32": "f4", "double": "f8", "float64": "f8",
}


def _np(x):
    if hasattr(x, "detach"):
        x = x.detach().cpu().numpy()
    return np.asarray(x)


def construct_list_of_attributes(n_dc, n_rest, n_scale=3, n_rot=4):
    """Attribute names in file order (scene/gaussian_model.py:263-275); n_dc / n_rest are the flattened widths."""
    names = ["x", "y", "z", "nx", "ny", "nz"]
    names += ["f_dc_%d" % i for i in range(n_dc)]
    names += ["f_rest_%d" % i for i in range(n_rest)]
    names.append("opacity")
    names += ["scale_%d" % i for i in range(n_scale)]
    names += ["rot_%d" % i for i in range(n_rot)]
    return names


def save_ply(path, xyz, features_dc, features_rest, opacity, scaling, rotation):
    """Write the model parameters as the reference does (scene/gaussian_model.py:277-296).

    xyz (P,3), features_dc (P,1,3), features_rest (P,K-1,3), opacity (P,1), scaling (P,3), rotation (P,4): the raw
    parameters `_xyz`, `_features_dc`, `_features_rest`, `_opacity`, `_scaling`, `_rotation`."""
    xyz = _np(xyz).astype(np.float32, copy=False)
    P = xyz.shape[0]
    if xyz.ndim != 2 or xyz.shape[1] != 3:
        raise ValueError("save_ply: xyz must have shape (P,3)")
    dc = _np(features_dc).astype(np.float32, copy=False)
    rest = _np(features_rest).astype(np.float32, copy=False)
    if dc.ndim != 3 or rest.ndim != 3 or dc.shape[0] != P or rest.shape[0] != P:
        raise ValueError("save_ply: features must have shape (P, coefficients, 3)")
    flat = lambda a: a.reshape(P, int(np.prod(a.shape[1:])))                 # explicit width: P may be 0
    f_dc = flat(np.ascontiguousarray(dc.transpose(0, 2, 1)))                 # channel-major, :281
    f_rest = flat(np.ascontiguousarray(rest.transpose(0, 2, 1)))             # :282
    op = flat(_np(opacity).astype(np.float32, copy=False))
    sc = flat(_np(scaling).astype(np.float32, copy=False))
    rot = flat(_np(rotation).astype(np.float32, copy=False))
    normals = np.zeros_like(xyz)
    attributes = np.concatenate((xyz, normals, f_dc, f_rest, op, sc, rot), axis=1)   # :293
    names = construct_list_of_attributes(f_dc.shape[1], f_rest.shape[1], sc.shape[1], rot.shape[1])
    assert attributes.shape[1] == len(names)
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)                                        # mkdir_p, :278
    header = ["ply", "format binary_little_endian 1.0", "element vertex %d" % P]
    header += ["property float %s" % n for n in names]
    header.append("end_header")
    with open(path, "wb") as f:
        f.write(("\n".join(header) + "\n").encode("ascii"))
        f.write(np.ascontiguousarray(attributes, dtype="<f4").tobytes())


def read_ply_vertices(path):
    """Parse a PLY file and return its `vertex` element as a numpy structured array (what `plydata.elements[0]` is for
    the reference's files) together with the property names in file order."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError("%s: not a PLY file" % path)
        fmt = None
        elements = []          # [name, count, [(prop, numpy code)]]
        while True:
            line = f.readline()
            if not line:
                raise ValueError("%s: header ends without end_header" % path)
            tok = line.decode("ascii", "replace").split()
            if not tok or tok[0] in ("comment", "obj_info"):
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                elements.append([tok[1], int(tok[2]), []])
            elif tok[0] == "property":
                if tok[1] == "list":
                    elements[-1][2].append((tok[-1], None))      # rejected below if it is in the element that is read
                else:
                    if tok[1] not in _PLY_TYPES:
                        raise ValueError("%s: unknown property type %r" % (path, tok[1]))
                    elements[-1][2].append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if fmt not in ("binary_little_endian", "binary_big_endian", "ascii"):
            raise ValueError("%s: unsupported PLY format %r" % (path, fmt))
        if not elements:
            raise ValueError("%s: no elements" % path)
        name, count, props = elements[0]       # the reference reads elements[0] (:313-316)
        if any(code is None for _, code in props):
            raise ValueError("%s: list properties are not supported in element %r" % (path, name))
        names = [p for p, _ in props]
        if fmt == "ascii":
            native = np.dtype([(p, "=" + c) for p, c in props])
            out = np.empty(count, dtype=native)
            rows = np.loadtxt(f, dtype=np.float64, max_rows=count, ndmin=2) if count else np.zeros((0, len(props)))
            if rows.shape != (count, len(props)):
                raise ValueError("%s: expected %d x %d ASCII values, found %s" % (path, count, len(props), rows.shape))
            for j, p in enumerate(names):
                out[p] = rows[:, j]
            return out, names
        order = "<" if fmt == "binary_little_endian" else ">"
        dt = np.dtype([(p, order + c) for p, c in props])
        raw = f.read(count * dt.itemsize)
        if len(raw) != count * dt.itemsize:
            raise ValueError("%s: truncated vertex data (%d of %d bytes)" % (path, len(raw), count * dt.itemsize))
        return np.frombuffer(raw, dtype=dt, count=count), names


def _sorted_names(names, prefix):
    sel = [n for n in names if n.startswith(prefix)]
    return sorted(sel, key=lambda n: int(n.split("_")[-1]))      # :328-329, :338-339, :344-345


def load_ply(path, max_sh_degree=3):
    """Read a model file the way GaussianModel.load_ply does (scene/gaussian_model.py:311-355) and return a dict of float32
    numpy arrays shaped as the model's parameters: xyz (P,3), features_dc (P,1,3), features_rest (P,K-1,3),
    opacity (P,1), scaling (P,3), rotation (P,4).  Raises AssertionError when the number of f_rest_* attributes does
    not match max_sh_degree, as the reference does (:330)."""
    v, names = read_ply_vertices(path)
    P = v.shape[0]
    xyz = np.stack((np.asarray(v["x"]), np.asarray(v["y"]), np.asarray(v["z"])), axis=1)
    opacities = np.asarray(v["opacity"])[..., np.newaxis]
    features_dc = np.zeros((P, 3, 1))
    for c in range(3):
        features_dc[:, c, 0] = np.asarray(v["f_dc_%d" % c])
    extra = _sorted_names(names, "f_rest_")
    assert len(extra) == 3 * (max_sh_degree + 1) ** 2 - 3
    features_extra = np.zeros((P, len(extra)))
    for i, n in enumerate(extra):
        features_extra[:, i] = np.asarray(v[n])
    features_extra = features_extra.reshape((P, 3, (max_sh_degree + 1) ** 2 - 1))   # (P, F, coefficients except DC), :335-336
    scale_names = _sorted_names(names, "scale_")
    scales = np.zeros((P, len(scale_names)))
    for i, n in enumerate(scale_names):
        scales[:, i] = np.asarray(v[n])
    rot_names = _sorted_names(names, "rot")
    rots = np.zeros((P, len(rot_names)))
    for i, n in enumerate(rot_names):
        rots[:, i] = np.asarray(v[n])
    f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
    return {
        "xyz": f32(xyz),
        "features_dc": f32(features_dc.transpose(0, 2, 1)),          # .transpose(1, 2).contiguous(), :352
        "features_rest": f32(features_extra.transpose(0, 2, 1)),     # :353
        "opacity": f32(opacities),
        "scaling": f32(scales),
        "rotation": f32(rots),
    }


def gaussians_from_ply(path, max_sh_degree=3, device="cuda"):
    """load_ply as torch tensors on `device` (the tensors GaussianModel.load_ply wraps in nn.Parameter, :350-356)."""
    import torch
    return {k: torch.from_numpy(a).to(device) for k, a in load_ply(path, max_sh_degree).items()}
