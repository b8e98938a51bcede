"""The streaming sweep's work decomposition (csrc/sweep_plan.h: which CTA sweeps which planes of which column) on
the CPU.  A plan decides nothing about the results as long as every plane of every column is swept exactly once and
the face segments are where the hand-over expects them -- which is what this checks, for the benchmark's shapes, for
awkward ones, for every plan part and for the knobs."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib(tmp_path_factory):
    so = tmp_path_factory.mktemp("plan") / "libplan.so"
    subprocess.run(["g++", "-std=c++17", "-O1", "-shared", "-fPIC", os.path.join(ROOT, "tests", "host_emul", "plan_check.cpp"),
                    "-o", str(so)], check=True)
    L = C.CDLL(str(so))
    L.plan_build.restype = C.c_int
    return L


def build(lib, Dl, H, nb, ty, part, lo=False, hi=False, cps=1, sm=148, edge=8, late=-1, zc=0, mode=2, fill=85, lastw=32, segcost=4):
    prm = (C.c_int * 16)(Dl, H, nb, ty, cps, part, sm, int(lo), int(hi), edge, late, zc, mode, fill, lastw, segcost)
    cap = 1 << 16
    segs = np.zeros((cap, 4), np.int32); first = np.zeros(cap, np.int32)
    ns, nc = C.c_int(0), C.c_int(0)
    built = lib.plan_build(prm, cap, cap, C.byref(ns), C.byref(nc), segs.ctypes.data_as(C.c_void_p), first.ctypes.data_as(C.c_void_p))
    assert ns.value <= cap and nc.value + 1 <= cap
    return built, segs[:ns.value].copy(), first[:nc.value + 1].copy()


def check(built, segs, first, Dl, H, nb, ty, lo, hi, edge):
    ncol = nb * ((H + ty - 1) // ty)
    assert first[0] == 0 and first[-1] == len(segs) and (np.diff(first) >= 0).all()
    # every plane of every column exactly once
    cover = np.zeros((ncol, Dl + 2), np.int32)
    for col, za, zb, mode in segs:
        assert 0 <= col < ncol and 1 <= za <= zb <= Dl, (col, za, zb)
        cover[col, za:zb + 1] += 1
    assert (cover[:, 1:Dl + 1] == 1).all(), "a plane is swept twice or not at all"
    face = segs[:, 3] >> 1
    if built in (4, 5, 6):
        for side, present in ((1, lo), (2, hi)):
            f = segs[face == side]
            assert len(f) == (ncol if present else 0)
            if present:
                assert sorted(f[:, 0]) == list(range(ncol))                       # one per column
                assert ((f[:, 2] - f[:, 1] + 1) >= 2).all()                       # two planes to hand over
                if side == 1:
                    assert (f[:, 1] == 1).all() and ((f[:, 3] & 1) == 0).all()    # starts at the face, streams up
                else:
                    assert (f[:, 2] == Dl).all() and ((f[:, 3] & 1) == 1).all()   # ends at the face, streams down
        if built in (5, 6):                                                       # face segments lead their CTA's list
            for k in range(len(first) - 1):
                m = face[first[k]:first[k + 1]] > 0
                assert not (np.diff(m.astype(int)) > 0).any()
    else:
        assert (face == 0).all()


SHAPES = [(256, 256, 1, 14), (256, 256, 4, 14), (256, 256, 3, 14), (64, 512, 4, 6), (128, 512, 2 * 4, 14), (33, 40, 1, 30),
          (40, 24, 3, 14), (257, 250, 2, 14), (36, 10, 2, 14), (512, 100, 1, 30)]


@pytest.mark.parametrize("Dl,H,nb,ty", SHAPES)
@pytest.mark.parametrize("part", [0, 4, 5, 6])
@pytest.mark.parametrize("faces", [(False, False), (True, False), (False, True), (True, True)])
def test_every_plane_once_and_faces_in_place(lib, Dl, H, nb, ty, part, faces):
    lo, hi = faces
    if part == 0 and (lo or hi):
        pytest.skip("part 0 has no face segments")
    if Dl < 32:
        pytest.skip("slabs thinner than 4 edges sweep whole columns")
    built, segs, first = build(lib, Dl, H, nb, ty, part, lo, hi)
    assert built in (part, 4)
    check(built, segs, first, Dl, H, nb, ty, lo, hi, 8)


@pytest.mark.parametrize("late", [-1, 0, 4, 12, 40])
@pytest.mark.parametrize("edge", [2, 4, 8])
def test_knobs_keep_the_cover(lib, late, edge):
    for part in (4, 6):
        built, segs, first = build(lib, 256, 256, 4, 14, part, True, True, edge=edge, late=late)
        check(built, segs, first, 256, 256, 4, 14, True, True, edge)
    for zc in (1, 3, 7):
        built, segs, first = build(lib, 40, 40, 2, 14, 0, zc=zc)
        check(built, segs, first, 40, 40, 2, 14, False, False, edge)
    for mode in (0, 1, 2):
        built, segs, first = build(lib, 256, 256, 1, 14, 0, mode=mode)
        check(built, segs, first, 256, 256, 1, 14, False, False, edge)
    for lastw, segcost in ((30, 5), (24, 4), (32, 7), (8, 1), (29, 15)):       # FLUID_SW_LASTW / FLUID_SW_SEGCOST
        for (Dl, H, nb, part, lo, hi) in ((256, 256, 4, 0, False, False), (256, 256, 1, 6, True, True), (100, 250, 3, 4, True, False)):
            built, segs, first = build(lib, Dl, H, nb, 14, part, lo, hi, edge=edge, late=late, lastw=lastw, segcost=segcost)
            check(built, segs, first, Dl, H, nb, 14, lo, hi, edge)


def test_benchmark_shapes_have_the_grids_the_traces_show(lib):
    """256^3 per GPU, <2,8> tiles (14 owned rows, 19 row tiles): the CTA counts of profiles/r02_slab_trace.txt."""
    assert len(build(lib, 256, 256, 4, 14, 0)[2]) - 1 == 148                      # one GPU, 4-field sweep
    assert len(build(lib, 256, 256, 1, 14, 0)[2]) - 1 == 148
    assert len(build(lib, 256, 256, 4, 14, 4, lo=True)[2]) - 1 == 224            # separate edge CTAs: 76 + 148
    assert len(build(lib, 256, 256, 1, 14, 4, lo=True)[2]) - 1 == 167            # 19 + 148
    built, segs, first = build(lib, 256, 256, 4, 14, 6, lo=True)                 # one wave
    assert built == 6 and len(first) - 1 == 148
    built, _, first = build(lib, 256, 256, 4, 14, 6, lo=True, hi=True)           # 152 face segments > 148 CTAs: falls back
    assert built == 4 and len(first) - 1 == 152 + 148
    built, _, first = build(lib, 256, 256, 1, 14, 6, lo=True, hi=True)
    assert built == 6 and len(first) - 1 == 148
