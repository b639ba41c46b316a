"""The C-ABI library loads without a GPU and exports every symbol include/b200sim.h declares."""
import ctypes as C
import os
import re

import pytest

from asimov_mjlab_b200 import abi, engine

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_declares_the_boundary():
  want = {"b2s_model_create", "b2s_model_destroy", "b2s_model_set_opt", "b2s_model_get_opt", "b2s_model_expand", "b2s_model_field",
          "b2s_data_create", "b2s_data_destroy", "b2s_data_field", "b2s_forward", "b2s_step", "b2s_step_n", "b2s_step_host",
          "b2s_last_error", "b2s_version", "b2s_built_arch", "b2s_launch_count", "b2s_profile_step"}
  assert want <= set(abi.EXPORTS)


def test_library_loads_and_exports_every_declared_symbol():
  L = engine.lib()
  for name in abi.EXPORTS:
    assert hasattr(L, name), name
  assert L.b2s_built_arch() == 100
  assert L.b2s_version() >= 100
  assert L.b2s_last_error() == b""


def test_no_torch_types_in_signatures():
  text = open(abi.HEADER).read()
  assert "torch" not in text.lower().replace("pytorch", "") or "at::" not in text
  assert 'extern "C"' in text
  assert not re.search(r"std::|at::|c10::", text)


def test_struct_mirror_matches_header_layout(g1):
  desc, keep = abi.make_model_desc(g1)
  assert desc.nq == 36 and desc.nv == 35 and desc.nu == 29 and desc.nbody == 32 and desc.ngeom == 69
  assert desc.opt_timestep == pytest.approx(0.005) and desc.opt_iterations == 10 and desc.opt_ls_iterations == 20
  assert desc.opt_integrator == abi.ENUMS["B2S_INT_IMPLICITFAST"] and desc.opt_solver == abi.ENUMS["B2S_SOL_NEWTON"]
  # pointer members are non-null and sized as the header's extent comments say
  dims = abi.model_dims(g1)
  for mem, ct, ptr, alen, ext in abi.MODEL_MEMBERS:
    if ptr:
      assert getattr(desc, mem), mem
      assert abi.eval_extent(ext, dims) >= 0


def test_model_create_without_gpu_fails_loudly_not_silently(g1):
  """No CPU fallback: on a machine without a CUDA device the product path raises."""
  import torch
  if torch.cuda.is_available():
    pytest.skip("GPU present")
  with pytest.raises(engine.BackendError):
    engine.put_model(g1, "cuda:0")
  with pytest.raises(engine.BackendError):
    engine.put_model(g1, "cpu")
  # and at the raw ABI level
  L = engine.lib()
  desc, keep = abi.make_model_desc(g1)
  h = C.c_void_p()
  rc = L.b2s_model_create(C.byref(desc), 0, C.byref(h))
  assert rc != 0 and L.b2s_last_error() != b""


def test_product_package_never_imports_the_oracle():
  pkg = os.path.join(ROOT, "asimov-mjlab_b200")
  for dirpath, _, files in os.walk(pkg):
    for f in files:
      if f.endswith((".py", ".cu", ".cuh", ".h")):
        text = open(os.path.join(dirpath, f)).read()
        assert not re.search(r"import\s+oracle|from\s+oracle|#include\s+\"[^\"]*oracle|liboracle|oracle/|orc_\w+\(", text), (dirpath, f)
