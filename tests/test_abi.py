"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports exactly the
symbols include/acnerf.h declares, the ctypes structs have the C layout, and the host
modules refuse CPU tensors (there is no CPU path)."""
import ctypes
import os
import re
import subprocess

import pytest
import torch

from conftest import ROOT


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "acnerf.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(acn_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from nerf_sys_b200 import _lib
    h = _lib.lib()
    syms = _header_symbols()
    assert len(syms) >= 35
    for s in syms:
        assert hasattr(h, s), f"{s} declared in acnerf.h but not exported"
        assert s in _lib.SIGNATURES, f"{s} has no ctypes signature"
    assert sorted(_lib.SIGNATURES) == syms
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = sorted(l.split()[-1] for l in out.splitlines() if " T " in l)
    assert exported == syms, "the shared object must export the C ABI and nothing else"
    assert h.acn_version() == 1
    assert h.acn_last_error() is not None


def test_struct_layout_matches_c():
    from nerf_sys_b200 import _lib
    assert ctypes.sizeof(_lib.AcnGridDesc) == 4 * 4 + 4 * 32
    assert ctypes.sizeof(_lib.AcnPoints) == 48  # x6, rays, t, S (+pad), P, t_T
    assert ctypes.sizeof(_lib.AcnPairs) == 32
    assert ctypes.sizeof(_lib.AcnNgpExpert) == 8 * 15 + 24
    assert ctypes.sizeof(_lib.AcnNgpExpertGrads) == 8 * 15
    src = "#include <stdio.h>\n#include \"acnerf.h\"\nint main(){printf(\"%zu %zu %zu %zu %zu\",sizeof(AcnGridDesc),sizeof(AcnPoints),sizeof(AcnPairs),sizeof(AcnNgpExpert),sizeof(AcnNgpExpertGrads));return 0;}"
    exe = "/tmp/acn_sizes"
    r = subprocess.run(["gcc", "-x", "c", "-", "-I", os.path.join(ROOT, "include"), "-o", exe], input=src, text=True, capture_output=True)
    assert r.returncode == 0, r.stderr  # the header is plain C
    sizes = list(map(int, subprocess.run([exe], capture_output=True, text=True).stdout.split()))
    assert sizes == [ctypes.sizeof(t) for t in (_lib.AcnGridDesc, _lib.AcnPoints, _lib.AcnPairs, _lib.AcnNgpExpert, _lib.AcnNgpExpertGrads)]


def test_invalid_arguments_return_status_not_crash():
    from nerf_sys_b200 import _lib
    h = _lib.lib()
    assert h.acn_ray_directions(0, 4, 1.0, 1.0, 0.0, 0.0, 1, None, None) == 1
    assert b"acn_ray_directions" in h.acn_last_error()
    g = _lib.AcnGridDesc()
    g.levels, g.features, g.log2_T, g.interp = 16, 4, 20, 0
    assert h.acn_hashgrid_fwd(None, 0, None, ctypes.byref(g), None, None) == 1
    assert b"features_per_level" in h.acn_last_error()
    assert h.acn_composite_fwd(None, None, None, 4, 1, 1.0, None, None, None, None, None) == 1


def test_cpu_tensors_are_refused():
    from nerf_sys_b200 import ops
    from nerf_sys_b200.models.encodings import HashGridEncoder
    with pytest.raises(RuntimeError, match="no CPU path"):
        ops.sh16(torch.randn(4, 3))
    enc = HashGridEncoder(levels=16, log2_hashmap_size=8)
    assert tuple(enc.hash_table.shape) == (16 * 256, 2)
    assert enc.level_resolutions.tolist() == [16, 23, 33, 48, 70, 101, 147, 212, 307, 445, 645, 933, 1351, 1955, 2830, 4095]
    with pytest.raises(RuntimeError, match="no CPU path"):
        enc(torch.rand(4, 3))


def test_state_dict_keys_match_reference_golden():
    """The drop-in container's state_dict is key- and shape-identical to the reference's
    (golden ngp_path.npz holds the reference model's own state_dict)."""
    from conftest import load_golden, state_from
    from nerf_sys_b200.models.inr import MetaContainer
    from nerf_sys_b200.nerfs.scene_box import SceneBox
    from oracle import acnerf_oracle as O
    z = load_golden("ngp_path")
    st = state_from(z)
    aabb, cent, boxes = O.grid_scene(1, 2)
    cfg = dict(O.HASH_CFG)
    cfg["log2_hashmap_size"] = int(z["log2_T"])
    m = MetaContainer(2, cent, aabb, "instant", boundary_margin=1.05, cluster_2d=True, use_bg_nerf=True, occ_conf={"use_occ": False},
                      expert_box_list=[SceneBox(b) for b in boxes], hidden=64, sigma_depth=2, color_depth=2, color_hidden=64,
                      dir_encoding="spherical", use_sigmoid_rgb=True, hash_enc_conf=cfg)
    own = m.state_dict()
    assert sorted(own) == sorted(st)
    for k in st:
        assert tuple(own[k].shape) == tuple(st[k].shape), k
    m.load_state_dict(st, strict=True)
    names = [n for n, _ in m.submodules[0].meta_named_parameters()]
    assert len(names) == 14 and "xyz_encoder.hash_table" not in names
    assert sorted(m.get_param_groups()) == ["background", "color", "encoding", "sigma"]
    assert float(m.submodules[0].__class__(occ_conf={}, scene_box=SceneBox(boxes[0]), color_depth=2, hash_enc_conf=cfg).sigma_head.bias) == -1.0
