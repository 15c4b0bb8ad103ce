"""The C-ABI library builds, loads, and exports every symbol include/particles_cuda.h declares.
No compute call is made here (no GPU in the CPU test tier)."""
import ctypes
import os
import re

import pytest

from particle_simulator_b200 import build, capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "particles_cuda.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ps_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_expected_surface():
    names = declared_symbols()
    for must in ("ps_create", "ps_destroy", "ps_upload", "ps_download", "ps_step", "ps_step_substeps", "ps_toggle_acc",
                 "ps_contact_pairs", "ps_last_error", "ps_integrate_bounce_f32"):
        assert must in names


def test_library_builds_and_exports_every_declared_symbol():
    path = build.build_lib()
    assert os.path.exists(path)
    L = ctypes.CDLL(path)
    for name in declared_symbols():
        assert hasattr(L, name), f"{name} declared in particles_cuda.h but not exported"


def test_ctypes_table_matches_the_header():
    assert sorted(capi.SYMBOLS) == declared_symbols()


def test_abi_version_and_status_strings():
    L = capi.load()
    assert L.ps_abi_version() == 1
    assert L.ps_status_string(0) == b"ok"
    assert b"CPU" not in L.ps_status_string(capi.PS_ERR_NO_DEVICE)
    assert ctypes.sizeof(capi.PsConfig) == 136  # must equal sizeof(ps_config); ps_create checks struct_size


def test_bad_arguments_are_rejected_without_touching_a_device():
    L = capi.load()
    h = ctypes.c_void_p()
    assert L.ps_create(None, ctypes.byref(h)) == capi.PS_ERR_BAD_ARG
    cfg = capi.PsConfig()
    cfg.struct_size = 4  # ABI mismatch
    assert L.ps_create(ctypes.byref(cfg), ctypes.byref(h)) == capi.PS_ERR_BAD_ARG
    assert b"struct_size" in L.ps_last_error(None)
    assert L.ps_step(None, 1) == capi.PS_ERR_BAD_ARG
    assert L.ps_upload(None, None, None) == capi.PS_ERR_BAD_ARG


def test_no_device_means_failure_not_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from particle_simulator_b200 import Simulation, scenes

    with pytest.raises(capi.PsError) as e:
        Simulation(scenes.default_scene())
    assert e.value.status == capi.PS_ERR_NO_DEVICE
