"""The C-ABI library loads and exports every symbol include/vga.h declares; without a CUDA device the
compute entry points fail loudly (there is no CPU fallback).  CPU only, no compute calls."""
import ctypes as C
import os
import re

import pytest

import vgaligner_b200 as vg
from conftest import ROOT, fixture_path


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "vga.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(vga_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported():
    lib = vg.gpu_lib()
    names = declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), "libvgaligner_gpu.so does not export %s" % n
    assert sorted(vg.GPU_SYMBOLS) == names


def test_integration_binding_lists_every_symbol():
    """The Rust extern block of INTEGRATION.md declares exactly the functions of include/vga.h."""
    text = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    assert sorted(set(re.findall(r"pub fn (vga_[a-z0-9_]+)", text))) == declared_symbols()


def header_structs():
    """field names, in order, of every `typedef struct { ... } Name;` of include/vga.h"""
    text = open(os.path.join(ROOT, "include", "vga.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    out = {}
    for m in re.finditer(r"typedef\s+struct\s*\w*\s*\{(.*?)\}\s*(\w+)\s*;", text, flags=re.S):
        fields = []
        for decl in m.group(1).split(";"):
            for part in decl.strip().split(","):
                if part.strip():
                    fields.append(re.findall(r"(\w+)\s*(?:\[\d+\])?$", part.strip())[0])
        out[m.group(2)] = fields
    return out


def test_struct_fields_agree_between_header_ctypes_and_rust():
    """Same field names in the same order in include/vga.h, the ctypes mirror and the #[repr(C)] structs of INTEGRATION.md
    (Rust spells `match` as `match_` and folds the trailing debug pointers into one array)."""
    structs = header_structs()
    assert set(structs) >= {"VgaIndexView", "VgaParams", "VgaResultsView", "VgaStats", "VgaGraphView"}
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    for name, fields in structs.items():
        assert [f for f, _ in getattr(vg, name)._fields_] == fields, name
        m = re.search(r"pub struct " + name + r"\s*\{(.*?)\n\}", doc, flags=re.S)
        assert m, name + " has no #[repr(C)] struct in INTEGRATION.md"
        body = re.sub(r"//[^\n]*", "", m.group(1))
        rust = []
        for fname, ftype in re.findall(r"pub (\w+)\s*:\s*([^,\n]+)", body):
            arr = re.match(r"\[\*const c_void;\s*(\d+)\]", ftype.strip())
            rust += [fname] * int(arr.group(1)) if arr else [fname.rstrip("_")]
        expect = [("dbg" if f.startswith("dbg_") else f) for f in fields]
        assert rust == expect, (name, rust, expect)


def test_no_oracle_in_product():
    """Nothing under the package or the CUDA library refers to oracle/."""
    pkg = os.path.join(ROOT, "rs-vgaligner_b200")
    for base, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(base, f), errors="ignore").read()
                assert "liboracle" not in src and "oracle/binding" not in src and "from oracle" not in src, f


def test_fails_loudly_without_device():
    lib = vg.gpu_lib()
    if lib.vga_device_count() > 0:
        pytest.skip("a CUDA device is present")
    assert lib.vga_init(0) == -2                      # VGA_ERR_NO_DEVICE
    assert b"no CPU fallback" in lib.vga_last_error()
    ix = vg.Index.build(fixture_path("test.gfa"), 4)
    with pytest.raises(vg.VgaError):
        ix.upload()
    with pytest.raises(vg.VgaError):
        vg.map_reads(ix, vg.read_seqs_from_file(fixture_path("single-read-test.fa")))


def test_params_defaults_match_reference():
    p = vg.VgaParams()
    vg.gpu_lib().vga_default_params(C.byref(p))
    # src/subcommands/map_main.rs:30-52,103 and abPOA's defaults
    assert (p.bandwidth, p.max_gap, p.min_anchors, p.best_n) == (50, 1000, 3, 1)
    assert (p.match, p.mismatch, p.o1, p.e1, p.o2, p.e2, p.wb) == (2, 4, 4, 2, 24, 1, 10)
    assert abs(p.wf - 0.01) < 1e-9
