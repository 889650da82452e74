"""The library's launch / report switches give the same results as the defaults:
DIBR_B200_REPORT=copy (list-size reports as copy nodes -- what a C caller handing pageable
memory gets), DIBR_B200_PDL=1 (programmatic dependent launch), DIBR_B200_RASTER_REC=0 (raster
lists of face ids instead of records) and DIBR_B200_GATHER_WORK=0 (the fused forward's gather
over every block of the screen instead of the rasterizer's work list).  All are read once
per process, so each combination runs the oracle-checked smoke in its own interpreter."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("env", [{"DIBR_B200_REPORT": "copy"}, {"DIBR_B200_PDL": "1"},
                                 {"DIBR_B200_REPORT": "copy", "DIBR_B200_PDL": "1"},
                                 {"DIBR_B200_WORKSPACE": "poll"}, {"DIBR_B200_WORKSPACE": "sync"},
                                 {"DIBR_B200_RASTER_REC": "0"}, {"DIBR_B200_GATHER_WORK": "0"}])
def test_smoke_under_switch(env):
    e = dict(os.environ)
    e.update(env)
    code = ("import __graft_entry__ as g, torch, workloads as wl, kaolin_windows_b200 as kal\n"
            "g.smoke()\n"
            "inp = wl.make_config('cfg1_pizza_256', device='cuda:0')\n"
            "a = kal.render.mesh.dibr_rasterization(256, 256, inp['face_vertices_z'], "
            "inp['face_vertices_image'], inp['face_features'], inp['face_normals_z'])\n"
            "b = kal.render.mesh.dibr_rasterization(256, 256, inp['face_vertices_z'], "
            "inp['face_vertices_image'], inp['face_features'], inp['face_normals_z'], "
            "cuda_graph=True)\n"
            "b = kal.render.mesh.dibr_rasterization(256, 256, inp['face_vertices_z'], "
            "inp['face_vertices_image'], inp['face_features'], inp['face_normals_z'], "
            "cuda_graph=True)\n"
            "assert all(torch.equal(x, y) for x, y in zip(a, b))\n"
            "print('modes ok')\n")
    r = subprocess.run([sys.executable, "-c", code], cwd=ROOT, env=e, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0 and "modes ok" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_lazy_workspace_queues_ahead_and_reports_overflow_late():
    """DIBR_B200_WORKSPACE=lazy (the default): after the first, verified call of a shape the
    eager API returns without waiting for the device; a list overflow of such a call surfaces
    at the next check as a RuntimeError, the capacity is raised and the re-run is right."""
    import torch
    import workloads as wl
    import kaolin_windows_b200 as kal
    from kaolin_windows_b200 import _workspace
    if _workspace.SYNC_MODE != "lazy":
        pytest.skip("lazy mode switched off")
    dev = "cuda:0"
    g = wl.fixture_geometry(batch_size=2)
    z, img, feat, nz = (g[k].to(dev) for k in ("face_vertices_z", "face_vertices_image",
                                               "face_uvs", "face_normals_z"))
    H = W = 257                                     # a shape no other test uses
    small = img * 0.02 + 0.37                       # every face well inside one bin
    want = kal.render.mesh.dibr_rasterization(H, W, z, img, feat, nz)   # never called lazily here
    _workspace._caps.pop((0, "render", 2, z.shape[1], H, W, 30), None)
    _workspace._lazy.pop((0, "render", 2, z.shape[1], H, W, 30), None)
    key = (0, "render", 2, z.shape[1], H, W, 30)
    kal.render.mesh.dibr_rasterization(H, W, z, small, feat, nz)        # calibrates on tiny lists
    need = _workspace._pinned_slot(key)             # (initial capacities are generous: shrink
    _workspace._caps[key][0] = int(need[0]) + 8     #  them to just what this geometry needs)
    _workspace._caps[key][1] = int(need[1]) + 8
    # queued without a host wait, several deep
    outs = [kal.render.mesh.dibr_rasterization(H, W, z, small, feat, nz) for _ in range(5)]
    assert len(_workspace._lazy[key].pending) >= 1
    _workspace.check_pending()
    assert all(torch.equal(o[2], outs[0][2]) for o in outs)
    # the spread-out geometry crosses bin borders and needs more entries: found at the next check
    kal.render.mesh.dibr_rasterization(H, W, z, img, feat, nz)
    with pytest.raises(RuntimeError, match="incomplete"):
        _workspace.check_pending()
    got = kal.render.mesh.dibr_rasterization(H, W, z, img, feat, nz)    # re-run: verified again
    _workspace.check_pending()
    assert all(torch.equal(a, b) for a, b in zip(got, want))
