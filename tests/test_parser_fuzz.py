"""Differential fuzz of the scene parser (csrc/host/scene_parser.cpp) against the REFERENCE
PROGRAM ITSELF (oracle/_ref/raytracer_ref = /root/reference/main.cpp compiled unmodified).

Seeded random scene files built from the reference's grammar and its quirks (trailing junk
after numbers, `mtlcolor` with 10 / 11 / 12 arguments, `bkgcolor` with and without eta, the four `f`
formats mixed per vertex, out-of-range vertex indices, comments, unknown commands, commands without
arguments, double blanks, missing required commands, unreadable textures ...).  For every file:

  reference wrote scene.ppm      -> our parser must accept the file, and parser -> CPU oracle -> our P3
                                    writer must give the reference's file byte for byte;
  reference printed "Error: ..." -> our parser must refuse with the same "Requires command" message;
  reference aborted (uncaught    -> our parser must refuse (RT_E_PARSE);
    exception, SIGABRT)
  reference died of SIGSEGV, or  -> undefined behaviour upstream: nothing to compare (ours must refuse an
    the file names a texture        unreadable texture).
    that cannot be opened

CPU only (tiny images); the fixture files are generated, nothing is read from /root/reference at run time.
"""
import os
import subprocess

import numpy as np
import pytest

from simple_raytracer_b200 import abi, ppm, scenegen
from simple_raytracer_b200.scene import Scene
from tests import oracle_lib as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "raytracer_ref")

pytestmark = pytest.mark.skipif(not os.path.exists(REF_BIN), reason="oracle/_ref/raytracer_ref not built")


def _num(rng, lo=-3.0, hi=3.0, junk=0.0):
    x = rng.uniform(lo, hi)
    style = rng.integers(0, 5)
    if style == 0:
        s = f"{x:.3f}"
    elif style == 1:
        s = f"{x:.2e}"
    elif style == 2:
        s = str(int(round(x)))
    elif style == 3:
        s = f"{x:+.4f}" if x >= 0 else f"{x:.4f}"
    else:
        s = f"{x:.1f}"
    if rng.random() < junk:
        s += rng.choice(["git", "f", "x7", "abc"])          # std::stof stops at the junk (A10.2)
    return s


def _make_scene(rng, workdir, with_faults):
    """-> scene text.  with_faults: sprinkle the malformed constructs the reference chokes on."""
    L = []
    req = {
        "eye": lambda: "eye " + " ".join(_num(rng, -1, 1) for _ in range(3)[:2]) + " " + _num(rng, 4, 7),
        "viewdir": lambda: f"viewdir {_num(rng, -0.3, 0.3)} {_num(rng, -0.3, 0.3)} -1",
        "updir": lambda: f"updir {_num(rng, -0.2, 0.2)} 1 {_num(rng, -0.2, 0.2)}",
        "hfov": lambda: "hfov " + _num(rng, 30, 100),
        "imsize": lambda: f"imsize {rng.integers(2, 9)} {rng.integers(2, 7)}",
        "bkgcolor": lambda: "bkgcolor " + " ".join(_num(rng, 0, 1) for _ in range(3)) + (" " + _num(rng, 1, 1.6) if rng.random() < 0.5 else ""),
    }
    order = list(req)
    rng.shuffle(order)
    for k in order:
        if with_faults and rng.random() < 0.04:
            continue                                         # a missing required command
        L.append(req[k]())
    nv = nvn = nvt = 0
    have_mtl = False
    n_cmds = rng.integers(3, 22)
    for _ in range(n_cmds):
        r = rng.random()
        junk = 0.08
        if r < 0.16:
            n_args = rng.choice([10, 10, 12, 12, 11] if not with_faults else [10, 12, 11, 9, 13])
            vals = [_num(rng, 0, 1, junk) for _ in range(6)] + [_num(rng, 0, 0.6), _num(rng, 0, 0.9), _num(rng, 0, 0.9), str(rng.integers(1, 60))]
            if n_args >= 11:
                vals.append(_num(rng, -0.3, 1.4))             # alpha, clamped by the parser
            if n_args >= 12:
                vals.append(_num(rng, 0.8, 1.8))
            if n_args == 13:
                vals.append("1")
            L.append("mtlcolor " + " ".join(vals[:n_args] if n_args <= 12 else vals))
            have_mtl = have_mtl or n_args in (10, 11, 12)
        elif r < 0.34:
            if not have_mtl and not with_faults:
                continue
            L.append("sphere " + " ".join(_num(rng, -2, 2, junk) for _ in range(3)) + " " + _num(rng, 0.3, 1.5, junk))
        elif r < 0.44:
            L.append("light " + " ".join(_num(rng, -4, 4) for _ in range(3)) + f" {rng.integers(0, 2)} " + " ".join(_num(rng, 0.2, 1) for _ in range(3)))
        elif r < 0.56:
            L.append("v " + " ".join(_num(rng, -2.5, 2.5, junk) for _ in range(3))); nv += 1
        elif r < 0.62:
            L.append("vn " + " ".join(_num(rng, -1, 1) for _ in range(3))); nvn += 1
        elif r < 0.68:
            L.append("vt " + " ".join(_num(rng, -0.2, 1.3) for _ in range(2))); nvt += 1
        elif r < 0.82:
            if nv < 1 or (not have_mtl and not with_faults):
                continue
            verts = []
            for _k in range(3):
                vi = rng.integers(1, nv + (3 if rng.random() < 0.1 else 1))        # sometimes out of range (A10.5)
                fmt = rng.integers(0, 4)
                ti = rng.integers(1, max(nvt, 1) + 1); ni = rng.integers(1, max(nvn, 1) + 1)
                verts.append([f"{vi}", f"{vi}/{ti}", f"{vi}//{ni}", f"{vi}/{ti}/{ni}"][fmt])
            L.append("f " + " ".join(verts))
        elif r < 0.86:
            L.append("# " + " ".join(str(x) for x in rng.integers(0, 99, 3)))
        elif r < 0.90:
            L.append(rng.choice(["foo 1 2 3", "spheres 0 0 0 1", "sphere", "v", "light", "texture"]))
        elif r < 0.95:
            if rng.random() < 0.7 or not with_faults:
                name = f"tex{rng.integers(0, 3)}.ppm"
                path = os.path.join(workdir, name)
                if not os.path.exists(path):
                    with open(path, "w") as f:
                        f.write(scenegen.p3_text(scenegen.make_texture("checker", 8, 4, int(rng.integers(0, 99)))))
                L.append("texture " + name)
            else:
                L.append("texture missing_file.ppm")
        elif with_faults:
            L.append(rng.choice(["sphere  0 0 0 1", " light 0 1 0 0 1 1 1", "imsize 1 5", "v 1 2", "hfov", "mtlcolor 1 1 1"]))
    return "\n".join(L) + "\n"


def _run_reference(workdir):
    r = subprocess.run([REF_BIN, "scene.txt"], cwd=workdir, capture_output=True, text=True, timeout=60)
    out_path = os.path.join(workdir, "scene.ppm")
    return r.returncode, r.stdout, (open(out_path, "rb").read() if os.path.exists(out_path) else None)


def _check_one(tmp, text):
    """-> one of 'same-file', 'both-refuse', 'both-message', 'ub'; asserts on a disagreement."""
    with open(os.path.join(tmp, "scene.txt"), "w") as f:
        f.write(text)
    rc, stdout, ref_ppm = _run_reference(tmp)
    ours_err = None
    parsed = None
    try:
        parsed = Scene.from_file(os.path.join(tmp, "scene.txt"), texture_dir=tmp)
    except abi.RtError as e:
        ours_err = e
    if "missing_file.ppm" in text:
        # utility.h:59-115: a texture file that cannot be opened leaves `width` / `height` uninitialised and
        # `new Texture(width, height)` runs on garbage: the reference aborts, or carries on, as the stack
        # happens to be.  Ours refuses, always.
        assert ours_err is not None
        return "ub"
    if rc == -11:
        return "ub"
    if rc != 0:                                            # SIGABRT: uncaught exception in the reference parser
        assert ours_err is not None, f"reference aborted (rc {rc}) but our parser accepted:\n{text}"
        return "both-refuse"
    if ref_ppm is None:                                    # the reference printed a message and returned 0
        assert ours_err is not None, f"reference refused ({stdout.strip()[-120:]}) but our parser accepted:\n{text}"
        if "Requires command" in stdout:
            want = [l for l in stdout.splitlines() if "Requires command" in l][-1].strip()
            assert want in ours_err.message, (want, ours_err.message, text)
        return "both-message"
    assert ours_err is None, f"reference rendered the file but our parser refused: {ours_err}\n{text}"
    raw, ids, _ = O.oracle_render(parsed)
    rgb, flags = O.raw_to_rgb8(raw, ids)
    ours = ppm.format_ppm(rgb.reshape(parsed.height, parsed.width, 3), flags.reshape(parsed.height, parsed.width), parsed.bkg)
    assert ours.encode() == ref_ppm, f"different files for:\n{text}"
    return "same-file"


@pytest.mark.parametrize("seed,with_faults", [(s, False) for s in range(8)] + [(100 + s, True) for s in range(16)])
def test_parser_agrees_with_the_reference_program_on_random_files(tmp_path, seed, with_faults):
    rng = np.random.default_rng(seed)
    tally = {}
    for k in range(30):
        d = tmp_path / f"case{k}"
        d.mkdir()
        text = _make_scene(rng, str(d), with_faults)
        kind = _check_one(str(d), text)
        tally[kind] = tally.get(kind, 0) + 1
    if not with_faults:
        assert tally.get("same-file", 0) >= 20, tally     # the well-formed generator mostly yields renderable files
