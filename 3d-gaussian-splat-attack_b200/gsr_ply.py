"""Gaussian-scene PLY I/O in the reference's on-disk layout, without the `plyfile` package.

Writer  scene/gaussian_model.py:377-411 (construct_list_of_attributes + save_ply): binary little-endian PLY, one
        `vertex` element, every property float32, in this order:
        x y z nx ny nz  f_dc_0..2  f_rest_0..(3*(M-1)-1)  opacity  scale_0..2  rot_0..3  obj_dc_0..(K-1)
        The SH blocks are stored CHANNEL-major: f_rest_{c*(M-1)+k} is coefficient k+1 of colour channel c
        (`transpose(1, 2).flatten(start_dim=1)`), likewise f_dc and obj_dc.
Reader  scene/gaussian_model.py:418-467 (load_ply): properties are looked up by NAME, f_rest/scale/rot sorted by
        their numeric suffix; a file without obj_dc_* loads with zero object features.

In memory the scene is a dict of float32 numpy arrays in the shapes the rasterizer's callers use
(scene/gaussian_model.py:97-121): xyz (P,3), features_dc (P,1,3), features_rest (P,M-1,3), opacity (P,1) raw
(pre-sigmoid), scaling (P,3) raw (log), rotation (P,4) raw (un-normalised), objects_dc (P,1,K).
`activated()` applies the activations render() applies (gaussian_renderer/__init__.py:53-81).
"""
import re

import numpy as np

__all__ = ["save_ply", "load_ply", "combine_splats", "activated", "property_names"]

_TYPES = {"float": "<f4", "float32": "<f4", "double": "<f8", "float64": "<f8", "uchar": "u1", "uint8": "u1",
          "char": "i1", "int8": "i1", "ushort": "<u2", "uint16": "<u2", "short": "<i2", "int16": "<i2",
          "uint": "<u4", "uint32": "<u4", "int": "<i4", "int32": "<i4"}


def property_names(n_rest, n_obj):
    """construct_list_of_attributes (scene/gaussian_model.py:377-392) for n_rest = 3*(M-1) and n_obj = K."""
    names = ["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(3)]
    names += [f"f_rest_{i}" for i in range(n_rest)] + ["opacity"]
    names += [f"scale_{i}" for i in range(3)] + [f"rot_{i}" for i in range(4)]
    names += [f"obj_dc_{i}" for i in range(n_obj)]
    return names


def _chan_major(a):
    """(P, coeffs, chan) -> (P, chan*coeffs): transpose(1, 2).flatten(start_dim=1)."""
    a = np.asarray(a, np.float32)
    return np.ascontiguousarray(np.transpose(a, (0, 2, 1))).reshape(a.shape[0], -1)


def save_ply(path, scene):
    xyz = np.asarray(scene["xyz"], np.float32)
    P = xyz.shape[0]
    obj = scene.get("objects_dc")
    if obj is None:
        obj = np.zeros((P, 1, 0), np.float32)
    cols = [xyz, np.zeros_like(xyz), _chan_major(scene["features_dc"]), _chan_major(scene["features_rest"]),
            np.asarray(scene["opacity"], np.float32).reshape(P, 1), np.asarray(scene["scaling"], np.float32).reshape(P, 3),
            np.asarray(scene["rotation"], np.float32).reshape(P, 4), _chan_major(obj)]
    table = np.ascontiguousarray(np.concatenate(cols, axis=1).astype("<f4"))
    names = property_names(cols[3].shape[1], cols[7].shape[1])
    assert table.shape[1] == len(names)
    header = "ply\nformat binary_little_endian 1.0\n" + f"element vertex {P}\n"
    header += "".join(f"property float {n}\n" for n in names) + "end_header\n"
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        f.write(table.tobytes())


def _read_vertex_table(path):
    """Structured array of the first element (`vertex`) of a binary-little-endian or ASCII PLY."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError(f"{path}: not a PLY file")
        fmt, count, props, in_first, n_elems = None, None, [], False, 0
        while True:
            line = f.readline()
            if not line:
                raise ValueError(f"{path}: unterminated PLY header")
            tok = line.decode("ascii", "replace").split()
            if not tok or tok[0] == "comment":
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                n_elems += 1
                in_first = n_elems == 1
                if in_first:
                    count = int(tok[2])
            elif tok[0] == "property" and in_first:
                if tok[1] == "list":
                    raise ValueError(f"{path}: list properties are not supported in the vertex element")
                if tok[1] not in _TYPES:
                    raise ValueError(f"{path}: unknown property type {tok[1]}")
                props.append((tok[2], _TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if count is None:
            raise ValueError(f"{path}: no element in header")
        dt = np.dtype(props)
        if fmt == "binary_little_endian":
            data = np.frombuffer(f.read(count * dt.itemsize), dtype=dt, count=count)
        elif fmt == "ascii":
            rows = [f.readline().split() for _ in range(count)]
            data = np.array([tuple(float(v) for v in r[: len(props)]) for r in rows], dtype=dt)
        else:
            raise ValueError(f"{path}: unsupported PLY format {fmt}")
    return data


def _numbered(names, prefix):
    pat = re.compile(re.escape(prefix) + r"(\d+)$")
    hits = sorted((int(m.group(1)), n) for n in names for m in [pat.match(n)] if m)
    return [n for _, n in hits]


def load_ply(path, max_sh_degree=3, num_objects=16):
    """load_ply (scene/gaussian_model.py:418-467): returns the scene dict described in the module docstring."""
    v = _read_vertex_table(path)
    names = v.dtype.names
    col = lambda n: np.asarray(v[n], np.float32)
    P = v.shape[0]
    xyz = np.stack([col("x"), col("y"), col("z")], axis=1)
    opacity = col("opacity")[:, None]
    dc = np.stack([col("f_dc_0"), col("f_dc_1"), col("f_dc_2")], axis=1)[:, None, :]        # (P,1,3)
    rest_names = _numbered(names, "f_rest_")
    n_rest = 3 * (max_sh_degree + 1) ** 2 - 3
    if len(rest_names) != n_rest:
        raise AssertionError(f"{path}: {len(rest_names)} f_rest_* properties, expected {n_rest}")  # :431
    rest = np.stack([col(n) for n in rest_names], axis=1) if rest_names else np.zeros((P, 0), np.float32)
    rest = rest.reshape(P, 3, (max_sh_degree + 1) ** 2 - 1).transpose(0, 2, 1)                # (P,M-1,3)
    scaling = np.stack([col(n) for n in _numbered(names, "scale_")], axis=1)
    rotation = np.stack([col(n) for n in _numbered(names, "rot_")], axis=1)
    objects = np.zeros((P, 1, num_objects), np.float32)
    if all(f"obj_dc_{i}" in names for i in range(num_objects)):
        objects[:, 0, :] = np.stack([col(f"obj_dc_{i}") for i in range(num_objects)], axis=1)
    c = np.ascontiguousarray
    return dict(xyz=c(xyz), features_dc=c(dc), features_rest=c(rest), opacity=c(opacity), scaling=c(scaling),
                rotation=c(rotation), objects_dc=objects)


def combine_splats(paths, max_sh_degree=3, num_objects=16, reference_masks=False):
    """combine_splats (scene/gaussian_model.py:469-556): concatenate several PLY scenes in the order given and
    return (scene, masks).  As in the reference, f_rest_* columns are zero-padded or truncated to 3*((D+1)^2-1)
    (a degree-0 splat can join a degree-3 scene), files need no obj_dc_* and the combined object features are all
    zero.  masks[k] is a bool vector over the COMBINED scene that is True on file k's Gaussians -- what the masked
    attacks index with.  The reference allocates mask k with only file k's length and sets [start, start+n_k) in
    it (:539-541), so every mask after the first comes out short and partly or wholly False; pass
    reference_masks=True to get exactly those."""
    if not paths:
        raise ValueError("No valid .ply files were loaded.")  # :545
    n_rest = 3 * ((max_sh_degree + 1) ** 2 - 1)
    parts, counts = [], []
    for path in paths:
        v = _read_vertex_table(path)
        names = v.dtype.names
        col = lambda n: np.asarray(v[n], np.float32)
        n = v.shape[0]
        rest_names = _numbered(names, "f_rest_")
        rest = np.stack([col(c) for c in rest_names], axis=1) if rest_names else np.zeros((n, 0), np.float32)
        if rest.shape[1] < n_rest:
            rest = np.hstack([rest, np.zeros((n, n_rest - rest.shape[1]), np.float32)])
        rest = rest[:, :n_rest].reshape(n, 3, n_rest // 3).transpose(0, 2, 1)
        parts.append(dict(
            xyz=np.stack([col("x"), col("y"), col("z")], axis=1),
            features_dc=np.stack([col("f_dc_0"), col("f_dc_1"), col("f_dc_2")], axis=1)[:, None, :],
            features_rest=rest, opacity=col("opacity")[:, None],
            scaling=np.stack([col(c) for c in _numbered(names, "scale_")], axis=1),
            rotation=np.stack([col(c) for c in _numbered(names, "rot_")], axis=1),
            objects_dc=np.zeros((n, 1, num_objects), np.float32)))
        counts.append(n)
    scene = {k: np.ascontiguousarray(np.concatenate([p[k] for p in parts], axis=0)) for k in parts[0]}
    total, start, masks = sum(counts), 0, []
    for n in counts:
        m = np.zeros(n if reference_masks else total, bool)
        m[start:start + n] = True
        masks.append(m)
        start += n
    return scene, masks


def activated(scene):
    """What render() feeds the rasterizer (gaussian_renderer/__init__.py:53-81): sigmoid opacity, exp scales,
    normalised quaternions, shs = cat(dc, rest) (P,M,3), sh_objs (P,1,K)."""
    q = np.asarray(scene["rotation"], np.float64)
    q = q / np.maximum(np.linalg.norm(q, axis=1, keepdims=True), 1e-12)
    f32 = np.float32
    return dict(means3D=np.asarray(scene["xyz"], f32),
                opacities=(1.0 / (1.0 + np.exp(-np.asarray(scene["opacity"], np.float64)))).astype(f32),
                scales=np.exp(np.asarray(scene["scaling"], np.float64)).astype(f32), rotations=q.astype(f32),
                shs=np.ascontiguousarray(np.concatenate([scene["features_dc"], scene["features_rest"]], axis=1), f32),
                sh_objs=np.asarray(scene.get("objects_dc", np.zeros((len(q), 1, 0))), f32))
