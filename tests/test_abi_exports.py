"""CPU-side checks of the C-ABI boundary: the library loads and exports every declared symbol."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "vdp_b200.h")


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vdp_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_expected_surface():
    syms = _declared_symbols()
    for must in ["vdp_btd_cholesky", "vdp_btd_inverse_band", "vdp_btd_solve", "vdp_gh_expect",
                 "vdp_naturals_to_ssm", "vdp_ssm_marginals", "vdp_cvi_kl_and_grad", "vdp_cvi_site_update"]:
        assert must in syms


def test_library_exports_every_declared_symbol():
    from vi_diffusion_processes_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    lib = _lib.load()
    for name in _declared_symbols():
        assert hasattr(lib, name), f"{name} declared in include/vdp_b200.h but not exported"
        assert name in _lib.SIGNATURES, f"{name} has no ctypes signature in _lib.py"
    assert lib.vdp_version() >= 100


@pytest.mark.parametrize("struct_name,cls_name", [("vdp_cvi_problem", "CviProblem"), ("vdp_vgp_problem", "VgpProblem"),
                                                  ("vdp_dist_sweep", "DistSweep")])
def test_struct_layout_matches_header_field_order(struct_name, cls_name):
    from vi_diffusion_processes_b200 import _lib
    CviProblem = getattr(_lib, cls_name)
    src = open(HEADER).read()
    body = src[src.index("typedef struct %s {" % struct_name):src.index("} %s;" % struct_name)]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for stmt in body.split("{", 1)[1].split(";"):
        for piece in stmt.split(","):
            m = re.search(r"\*?\s*([A-Za-z_][A-Za-z0-9_]*)\s*(\[\d+\])?\s*$", piece.strip())
            if m and piece.strip():
                names.append(m.group(1))
    assert names == [f[0] for f in CviProblem._fields_]


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from vi_diffusion_processes_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(_lib.VdpError):
        _lib.load()
