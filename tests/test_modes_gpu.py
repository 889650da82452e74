"""The library's launch / report switches give the same results as the defaults:
DIBR_B200_REPORT=copy (list-size reports as copy nodes -- what a C caller handing pageable
memory gets) and DIBR_B200_PDL=1 (programmatic dependent launch).  Both are read once per
process, so each combination runs the oracle-checked smoke in its own interpreter."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("env", [{"DIBR_B200_REPORT": "copy"}, {"DIBR_B200_PDL": "1"},
                                 {"DIBR_B200_REPORT": "copy", "DIBR_B200_PDL": "1"}])
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
