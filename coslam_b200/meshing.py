"""Inference path of mesh extraction: the dense-lattice SDF query and the per-vertex colour
query of `extract_mesh` (utils.py:80-201), without the 65 536-point chunk loop and its per-chunk host round trips
(`fn(flat, i, i + chunk).cpu().data.numpy()`, utils.py:129; 240 chunks at room0's voxel_final 0.02).

The lattice coordinates are the reference's own `torch.linspace` values (created on the CPU, utils.py:46-48); the grid is
assembled and normalised on the device with the same fp32 operations, so every query point -- and, because the per-point
kernels do not mix rows, every SDF / colour value -- is bit-identical to what the chunked loop produces
(tests/test_parity_gpu.py::test_mesh_lattice_query_matches_chunked_loop).  Marching cubes itself (a C++ extension of the
reference) and the trimesh export stay out of scope.
"""
import torch


def get_voxels(x_max, x_min, y_max, y_min, z_max, z_min, voxel_size=None, resolution=None):
    """getVoxels, utils.py:29-54."""
    x_max, x_min, y_max, y_min, z_max, z_min = (float(v) for v in (x_max, x_min, y_max, y_min, z_max, z_min))
    if voxel_size is not None:
        nx = round((x_max - x_min) / voxel_size + 0.0005)
        ny = round((y_max - y_min) / voxel_size + 0.0005)
        nz = round((z_max - z_min) / voxel_size + 0.0005)
        return torch.linspace(x_min, x_max, nx + 1), torch.linspace(y_min, y_max, ny + 1), torch.linspace(z_min, z_max, nz + 1)
    return torch.linspace(x_min, x_max, resolution), torch.linspace(y_min, y_max, resolution), torch.linspace(z_min, z_max, resolution)


@torch.no_grad()
def query_sdf_lattice(model, config, bounding_box, marching_cube_bound=None, voxel_size=None, resolution=None, chunk=1 << 22):
    """The lattice part of extract_mesh (utils.py:96-131): returns (sdf [Nx,Ny,Nz] fp32 on the device, (tx, ty, tz) CPU
    linspaces).  `model.query_sdf` is called on at most `chunk` points at a time, device to device."""
    dev = bounding_box.device
    if marching_cube_bound is None:
        marching_cube_bound = bounding_box
    x_min, y_min, z_min = marching_cube_bound[:, 0]
    x_max, y_max, z_max = marching_cube_bound[:, 1]
    tx, ty, tz = get_voxels(x_max, x_min, y_max, y_min, z_max, z_min, voxel_size, resolution)
    pts = torch.stack(torch.meshgrid(tx.to(dev), ty.to(dev), tz.to(dev), indexing='ij'), -1).to(torch.float32)
    shape = pts.shape[:-1]
    flat = pts.reshape(-1, 3)
    if config['grid']['tcnn_encoding']:                            # utils.py:112-115, same two fp32 operations
        flat = (flat - bounding_box[:, 0]) / (bounding_box[:, 1] - bounding_box[:, 0])
    out = torch.empty(flat.shape[0], dtype=torch.float32, device=dev)
    for i in range(0, flat.shape[0], chunk):
        out[i:i + chunk] = model.query_sdf(flat[i:i + chunk, None, :]).reshape(-1)
    return out.reshape(shape), (tx, ty, tz)


@torch.no_grad()
def query_vertex_colors(model, config, bounding_box, vertices, chunk=1 << 22):
    """Per-vertex colour of extract_mesh with `mesh.render_color: False` (utils.py:160-176): vertices [V,3] in world units
    (numpy or tensor) -> rgb [V,3] on the device through `model.query_color`."""
    v = torch.as_tensor(vertices).to(bounding_box)
    if config['grid']['tcnn_encoding']:
        v = (v - bounding_box[:, 0]) / (bounding_box[:, 1] - bounding_box[:, 0])
    out = torch.empty((v.shape[0], 3), dtype=torch.float32, device=v.device)
    for i in range(0, v.shape[0], chunk):
        out[i:i + chunk] = model.query_color(v[i:i + chunk, None, :]).reshape(-1, 3)
    return out
