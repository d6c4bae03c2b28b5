"""tools/sil_bench.py — time the soft-silhouette rasteriser (forward, forward+backward) at FIND's scale:
10 242-vertex template, 256 x 256, feet x views images per call.  Prints one JSON line."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from find_b200 import synthetic as syn                      # noqa: E402
from find_b200.mesh import Meshes                           # noqa: E402
from find_b200.renderer import FootRenderer                 # noqa: E402


def timed(fn, n=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def main():
    dev = torch.device("cuda")
    feet, size = int(os.environ.get("FEET", 8)), int(os.environ.get("SIZE", 256))
    tv, tf = syn.make_template(10242)
    verts = (tv[None] + 0.002 * torch.randn(feet, tv.shape[0], 3)).to(dev)
    r = FootRenderer(size, device=dev)
    R, T = r.view_from(["topdown", "side1", "side2", "45"])
    mesh_faces = tf.to(dev)
    out = {}
    with torch.no_grad():
        out["fwd_ms"] = timed(lambda: r(Meshes(verts, mesh_faces), R, T))
    colours = torch.rand(feet, tv.shape[0], 3, device=dev)
    with torch.no_grad():                           # Phong image (hard rasterisation + vertex normals + shading), forward only
        out["image_fwd_ms"] = timed(lambda: r(Meshes(verts, mesh_faces, textures=colours), R, T, return_images=True, return_mask=False))
    vg = verts.clone().requires_grad_(True)
    target = torch.rand(feet, 4, size, size, device=dev)

    def fb():
        vg.grad = None
        m = r(Meshes(vg, mesh_faces), R, T)["mask"]
        ((m - target) ** 2).mean().backward()
    out["fwd_bwd_ms"] = timed(fb)
    images = feet * 4
    out.update(images=images, image_size=size, faces=int(tf.shape[0]), mpix_per_s_fwd=images * size * size / out["fwd_ms"] / 1e3,
               images_per_s_fwd_bwd=images / out["fwd_bwd_ms"] * 1e3)
    if os.environ.get("CPU", "1") != "0":           # CPU port (oracle/raster_oracle.py, row-blocked) on ONE image of the same workload
        import time
        from oracle import raster_oracle as ro
        ndc = r.project(verts[:1].cpu(), R[:1], T[:1])
        t0 = time.perf_counter()
        want = ro.silhouette_blocked(ndc[0], tf, size, size)
        cpu_s = time.perf_counter() - t0
        with torch.no_grad():
            got = r(Meshes(verts[:1], mesh_faces), R[:1], T[:1])["mask"][0, 0].cpu()
        err = (got - want).abs()
        out["cpu_port"] = dict(s_per_image_fwd=cpu_s, images_per_s_fwd=1.0 / cpu_s, cores=torch.get_num_threads(), kind="port",
                               sample="1 of the 32 images, forward only, oracle/raster_oracle.py::silhouette_blocked (torch CPU)")
        out["parity_full_image"] = dict(max_abs_err=float(err.max()), mean_abs_err=float(err.mean()))
    print(json.dumps(out))


if __name__ == "__main__":
    main()
