"""Regenerate the compiled model blobs under asimov-mjlab_b200/assets/ from the
reference checkout's MJCF files (run in the build container; the GPU box has no
/root/reference).  Usage: python tools/export_models.py [/root/reference]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from asimov_mjlab_b200 import scenes  # noqa: E402

ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
zoo = os.path.join(ref, "src/mjlab/asset_zoo/robots")
m = scenes.build_g1_flat(os.path.join(zoo, "unitree_g1/xmls/g1.xml"))
scenes.save_model(m, os.path.join(scenes.ASSET_DIR, "g1_flat.npz"))
print("g1_flat: nq", m.nq, "nv", m.nv, "nu", m.nu, "nbody", m.nbody, "ngeom", m.ngeom,
      "nsite", m.nsite, "nsensordata", m.nsensordata, "meaninertia", m.stat.meaninertia)

t = scenes.build_g1_tracking(os.path.join(zoo, "unitree_g1/xmls/g1.xml"))
scenes.save_model(t, os.path.join(scenes.ASSET_DIR, "g1_tracking.npz"))
print("g1_tracking: nsensordata", t.nsensordata, "sensors", t.names["sensor"])

for terr in ("plane", "rough", "rough_hf"):
  g = scenes.build_go1(os.path.join(zoo, "unitree_go1/xmls/go1.xml"), terrain=terr)
  name = {"plane": "go1_flat", "rough": "go1_rough", "rough_hf": "go1_rough_hf"}[terr]
  scenes.save_model(g, os.path.join(scenes.ASSET_DIR, name + ".npz"))
  print(name, ": nq", g.nq, "nv", g.nv, "nu", g.nu, "nbody", g.nbody, "ngeom", g.ngeom,
        "nsite", g.nsite, "nsensordata", g.nsensordata, "meaninertia", g.stat.meaninertia)

for terr in ("plane", "rough"):
  a = scenes.build_asimov_toe(os.path.join(zoo, "asimov/xmls/asimov_toe.xml"), terrain=terr)
  name = "asimov_toe_flat" if terr == "plane" else "asimov_toe_rough"
  scenes.save_model(a, os.path.join(scenes.ASSET_DIR, name + ".npz"))
  print(name, ": nq", a.nq, "nv", a.nv, "nu", a.nu, "nbody", a.nbody, "ngeom", a.ngeom, "ntendon", a.ntendon,
        "nmesh", a.nmesh, "nmeshvert", a.nmeshvert, "nsensordata", a.nsensordata, "meaninertia", a.stat.meaninertia)

a = scenes.build_asimov_toe(os.path.join(zoo, "asimov/xmls/asimov_toe.xml"), terrain="plane", collision="full")
scenes.save_model(a, os.path.join(scenes.ASSET_DIR, "asimov_toe_full_flat.npz"))
print("asimov_toe_full_flat: ngeom", a.ngeom, "nmesh", a.nmesh, "nmeshvert", a.nmeshvert)

for col in ("feet_only", "full"):
  a = scenes.build_asimov(os.path.join(zoo, "asimov/xmls/asimov.xml"), collision=col)
  name = "asimov_flat" if col == "feet_only" else "asimov_full_flat"
  scenes.save_model(a, os.path.join(scenes.ASSET_DIR, name + ".npz"))
  print(name, ": nq", a.nq, "nv", a.nv, "nu", a.nu, "nbody", a.nbody, "ngeom", a.ngeom, "nmesh", a.nmesh, "nmeshvert", a.nmeshvert,
        "nsensordata", a.nsensordata, "npair-capable geoms", int(((a.geom_contype | a.geom_conaffinity) != 0).sum()))
