"""PLY files of mesh-anchored Gaussians: the `point_cloud.ply` the reference writes with
`plyfile` (`scene_reconstruction/gaussian_mesh.py:433-465`, `gaussian_model.py:181-213`): ONE element `vertex`, every
property a little-endian float32, in the order x y z nx ny nz f_dc_* f_rest_* opacity scale_* rot_* b1 b2 b3 o id
(the face id is stored as a float, like the reference does).  `plyfile` is not in the image; the format is small enough
to read and write with numpy: header lines as plyfile emits them, then N x K packed float32."""
import os

import numpy as np


def gaussian_attribute_names(n_dc: int, n_rest: int, n_scale: int = 3, n_rot: int = 4, mesh_anchored: bool = True):
    """`construct_list_of_attributes` of GaussianModel (gaussian_model.py:181-193) + MultiGaussianMesh's extension
    (gaussian_mesh.py:433-436)."""
    names = ["x", "y", "z", "nx", "ny", "nz"]
    names += [f"f_dc_{i}" for i in range(n_dc)]
    names += [f"f_rest_{i}" for i in range(n_rest)]
    names += ["opacity"]
    names += [f"scale_{i}" for i in range(n_scale)]
    names += [f"rot_{i}" for i in range(n_rot)]
    if mesh_anchored:
        names += ["b1", "b2", "b3", "o", "id"]
    return names


def write_vertex_ply(path: str, names, table: np.ndarray):
    """table: [N, len(names)] -> binary_little_endian PLY with one `vertex` element of float properties."""
    table = np.ascontiguousarray(table, dtype="<f4")
    if table.ndim != 2 or table.shape[1] != len(names):
        raise ValueError(f"table must be [N, {len(names)}]")
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    header = ["ply", "format binary_little_endian 1.0", f"element vertex {table.shape[0]}"]
    header += [f"property float {n}" for n in names]
    header += ["end_header"]
    with open(path, "wb") as f:
        f.write(("\n".join(header) + "\n").encode("ascii"))
        f.write(table.tobytes())


_PLY_TYPES = {"float": "f4", "float32": "f4", "double": "f8", "float64": "f8", "int": "i4", "int32": "i4",
              "uint": "u4", "uint32": "u4", "short": "i2", "int16": "i2", "ushort": "u2", "uint16": "u2",
              "char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1"}


def read_vertex_ply(path: str):
    """-> dict property name -> float64/ints numpy column of the FIRST element (binary little/big endian or ascii; scalar
    properties only, which is all the reference writes)."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError("not a PLY file")
        fmt, count, props, in_first, seen_first = None, None, [], False, False
        while True:
            line = f.readline()
            if not line:
                raise ValueError("PLY header without end_header")
            tok = line.decode("ascii").split()
            if not tok or tok[0] == "comment" or tok[0] == "obj_info":
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                in_first = not seen_first
                if in_first:
                    count, seen_first = int(tok[2]), True
            elif tok[0] == "property" and in_first:
                if tok[1] == "list":
                    raise ValueError("list properties are not supported")
                props.append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if fmt is None or count is None:
            raise ValueError("PLY header without format / element")
        if fmt == "ascii":
            rows = [f.readline().split() for _ in range(count)]
            return {n: np.asarray([r[i] for r in rows], dtype=t) for i, (n, t) in enumerate(props)}
        end = "<" if fmt == "binary_little_endian" else ">"
        dt = np.dtype([(n, end + t) for n, t in props])
        data = np.frombuffer(f.read(count * dt.itemsize), dtype=dt, count=count)
        return {n: np.asarray(data[n]) for n, _ in props}
