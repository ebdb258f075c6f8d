"""CPU-only tests of the host front-end (parser, P3 reader / writer) and of the
C-ABI libraries' export tables.  No compute call touches a GPU here."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from simple_raytracer_b200 import abi, ppm, scenegen
from simple_raytracer_b200.scene import Scene
from tests import common
from tests import oracle_lib as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "raytracer_ref")


def _declared(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported():
    rt, host = abi.load_rt(), abi.load_host()
    names_rt, names_host = _declared("rt.h"), [n for n in _declared("rt_host.h") if n.startswith("rt_host_")]
    assert set(names_rt) == set(abi.RT_SYMBOLS)
    assert set(names_host) == set(abi.RT_HOST_SYMBOLS)
    for n in names_rt:
        assert getattr(rt, n) is not None
    for n in names_host:
        assert getattr(host, n) is not None
    assert rt.rt_abi_version() == 4


def test_struct_layouts_match_the_header():
    # sizes implied by include/rt.h on LP64
    assert C.sizeof(abi.RtMaterial) == 48 and C.sizeof(abi.RtSphere) == 28
    assert C.sizeof(abi.RtFace) == 112 and C.sizeof(abi.RtLight) == 28 and C.sizeof(abi.RtTexture) == 16
    assert C.sizeof(abi.RtRenderParams) == 40
    assert C.sizeof(abi.RtSceneDesc) == 8 + 12 * 3 + 4 + 12 + 4 + 5 * 16 + 8


def _write(tmp_path, text, name="s.txt"):
    p = tmp_path / name
    p.write_text(text)
    return str(p)


BASE = "eye 0 0 5\nviewdir 0 0 -1\nupdir 0 1 0\nhfov 60\nimsize 8 6\nbkgcolor 0.1 0.2 0.3\n"


def test_parser_reads_the_example_scenes():
    s = common.example_scene("basic_geometry_tests/four_spheres")
    assert (s.width, s.height) == (512, 512) and len(s.spheres) == 4 and len(s.lights) == 1
    assert s.spheres["radius"][3] == 1.0                       # "sphere 0 0 -6 1git": stof accepts trailing junk (A10.2)
    assert s.bkg_eta == 0.0                                     # bkgcolor without eta (A10.3)
    assert list(s.spheres["id"]) == [1, 2, 3, 4]
    t1 = common.example_scene("reflection_transparency_tests/Test1")
    assert t1.bkg_eta == 1.0 and len(t1.faces) == 2 and len(t1.spheres) == 4
    assert list(t1.faces["id"]) == [5, 6]                      # ids count spheres and faces together (main.cpp:329,476)
    assert np.allclose(t1.materials["opacity"][:2], [0.2, 1.0]) and t1.materials["eta"][1] == 0.0
    h = common.example_scene("showcases/house")
    assert len(h.textures) == 4 and h.faces["texture"].max() == 2 and h.spheres["texture"][0] == 3
    assert h.faces["vt"].max() == 4.0                          # UVs beyond 1 are kept; the renderer clamps (A8.5)
    ep = common.example_scene("showcases/earth_pyramid")
    assert (ep.faces["vt"][2:] == 0).all() or (ep.faces["vt"] == 0).any()   # missing vt index reads as zeros (A10.5)


def test_parser_quirks(tmp_path):
    # mtlcolor alpha/eta only with exactly 12 arguments (A10.3)
    s = Scene.from_file(_write(tmp_path, BASE + "mtlcolor 1 0 0 1 1 1 0.1 0.2 0.3 4 0.5\nsphere 0 0 0 1\n"))
    assert s.materials["opacity"][0] == 1.0 and s.materials["eta"][0] == 1.0
    s = Scene.from_file(_write(tmp_path, BASE + "mtlcolor 1 0 0 1 1 1 0.1 0.2 0.3 4 7.5 1.5\nsphere 0 0 0 1\n"))
    assert s.materials["opacity"][0] == 1.0 and s.materials["eta"][0] == 1.5       # alpha clamped to [0,1]
    # comment lines and unknown commands are skipped; a command without arguments too
    s = Scene.from_file(_write(tmp_path, BASE + "# a comment here\nfoo 1 2 3\nsphere\nlight 0 -1 0 0 1 1 1\n"))
    assert len(s.spheres) == 0 and len(s.lights) == 1 and s.lights["w"][0] == 0
    # face format sniffing per vertex; the last vertex decides smooth shading (A10.6)
    txt = BASE + "mtlcolor 1 1 1 1 1 1 0 1 0 1\nv 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nvt 0.5 0.25\n"
    s = Scene.from_file(_write(tmp_path, txt + "f 1//1 2//1 3\nf 1 2 3//1\nf 1/1 2/1 3/1\nf 1/1/1 2/1/1 3/1/1\n"))
    assert list(s.faces["smooth"]) == [0, 1, 0, 1]
    assert s.faces["texture"].tolist() == [-1, -1, -1, -1]      # has_texture comes from the parser state (A8.6)
    assert s.faces["vn"][0][2].tolist() == [0, 0, 0] and s.faces["vn"][0][0].tolist() == [0, 0, 1]
    assert s.faces["vt"][2][1].tolist() == [0.5, 0.25]
    # a face that references a missing vertex silently gets zeros (A10.5)
    s = Scene.from_file(_write(tmp_path, txt + "f 1 2 9\n"))
    assert s.faces["v"][0][2].tolist() == [0, 0, 0]


@pytest.mark.parametrize("text,needle", [
    (BASE.replace("eye 0 0 5", "eye  0 0 5"), "empty token"),                    # double space aborts (A10.1)
    (BASE + " sphere 0 0 0 1\n", "empty token"),                                  # leading space aborts
    (BASE.replace("imsize 8 6", "imsize 1 6"), "Invalid image dimensions"),       # main.cpp:242
    (BASE + "sphere 0 0 0 1\n", "Must define a 'mtlcolor'"),                      # main.cpp:360
    (BASE.replace("hfov 60\n", ""), "Requires command 'hfov'"),                   # main.cpp:594
    (BASE.replace("bkgcolor 0.1 0.2 0.3\n", ""), "Requires command 'bkgcolor'"),
    (BASE + "mtlcolor 1 1 1\n", "material"),
    (BASE + "mtlcolor 1 1 1 1 1 1 0 1 0 1\ntexture nope.ppm\n", "texture"),
])
def test_parser_errors(tmp_path, text, needle):
    with pytest.raises(abi.RtError) as e:
        Scene.from_file(_write(tmp_path, text))
    assert e.value.code == abi.RT_E_PARSE and needle in e.value.message


def _parse_with(path, threads, chunk_bytes):
    """-> ('ok', every array of the scene as bytes) or ('err', code, message)"""
    old = {k: os.environ.get(k) for k in ("RT_HOST_THREADS", "RT_HOST_CHUNK_BYTES")}
    os.environ["RT_HOST_THREADS"] = str(threads)
    os.environ["RT_HOST_CHUNK_BYTES"] = str(chunk_bytes)
    try:
        s = Scene.from_file(path)
        cam = np.array([s.width, s.height, *s.eye, *s.viewdir, *s.updir, s.hfov_deg, *s.bkg, s.bkg_eta], dtype=np.float64)
        return ("ok", cam.tobytes(), s.materials.tobytes(), s.spheres.tobytes(), s.faces.tobytes(), s.lights.tobytes(),
                tuple(t.tobytes() for t in s.textures))
    except abi.RtError as e:
        return ("err", e.code, e.message)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def test_chunked_parallel_parse_equals_the_sequential_one(tmp_path):
    """The parser cuts the file into chunks parsed by several threads and links them afterwards
    (scene_parser.cpp): whatever the cut points, the scene — object ids, the material / texture an
    object captured, forward and missing vertex references, vertex counts — and the FIRST error
    must be those of the one-chunk (= line by line) parse."""
    rng = np.random.default_rng(5)
    tex = tmp_path / "t.ppm"
    tex.write_text(scenegen.p3_text(scenegen.make_texture("checker", 4, 4, 1)))
    texts = []
    # generated scenes: state restated per object (to_text) and a soup with shared state
    sc = scenegen.coverage_scene(16, 12)
    paths = []
    for i, t in enumerate(sc.textures):
        q = tmp_path / f"c{i}.ppm"
        q.write_text(scenegen.p3_text(t))
        paths.append(str(q))
    texts.append(sc.to_text(paths))
    soup = scenegen.random_soup(300, 20, 9)
    spaths = []
    for i, t in enumerate(soup.textures):
        q = tmp_path / f"s{i}.ppm"
        q.write_text(scenegen.p3_text(t))
        spaths.append(str(q))
    texts.append(soup.to_text(spaths))
    # hand-made: state carried across many lines, forward / missing / negative references, every face form
    lines = [BASE.strip(), "light 0 -1 0 0 1 1 1", "bkgcolor 0.3 0.2 0.1 1.5", "bkgcolor 0.5 0.5 0.5"]
    nv = nvn = nvt = 0
    for k in range(400):
        r = rng.integers(0, 10)
        if r == 0:
            lines.append("mtlcolor " + " ".join(f"{x:.3f}" for x in rng.uniform(0, 1, 10)) + (" 0.5 1.3" if k % 3 else ""))
        elif r == 1:
            lines.append(f"texture {tex}")
        elif r == 2 and k > 20:
            lines.append("sphere " + " ".join(f"{x:.2f}" for x in rng.uniform(-3, 3, 4)))
        elif r in (3, 4):
            lines.append("v " + " ".join(f"{x:.4f}" for x in rng.uniform(-3, 3, 3))); nv += 1
        elif r == 5:
            lines.append("vn " + " ".join(f"{x:.4f}" for x in rng.uniform(-1, 1, 3))); nvn += 1
        elif r == 6:
            lines.append("vt " + " ".join(f"{x:.4f}" for x in rng.uniform(0, 1, 2))); nvt += 1
        elif k > 20:
            def ref(n):
                return int(rng.integers(-1, n + 4))          # includes 0, -1 and not-yet-defined keys
            form = rng.integers(0, 4)
            spec = []
            for _ in range(3):
                a, b, c = ref(nv), ref(nvt), ref(nvn)
                spec.append([f"{a}", f"{a}/{b}", f"{a}//{c}", f"{a}/{b}/{c}"][(form + rng.integers(0, 2)) % 4])
            lines.append("f " + " ".join(spec))
    good = "\n".join(lines[:6] + ["mtlcolor 1 1 1 1 1 1 0 1 0 1"] + lines[6:]) + "\n"
    texts.append(good)
    texts.append(good.rstrip("\n"))                                  # no newline at the end
    texts.append(good.replace("\n", "\r\n"))                         # CRLF: the \r stays in the last token
    # failing files: the first error in file order wins, wherever the cuts fall
    texts.append("\n".join(lines) + "\n")                             # objects before any mtlcolor (maybe)
    bad = good.split("\n")
    for at, what in ((50, "v 1 2"), (120, "f 1 2"), (200, "sphere 1 2 x 4"), (260, "light  0 0 0 1 1 1 1"),
                     (300, "texture missing.ppm"), (10, "imsize 1 1")):
        b2 = list(bad)
        b2.insert(at, what)
        b2.insert(at + 40, "mtlcolor 1 1")                           # a later error must not win
        texts.append("\n".join(b2))
    # state errors: an object before any mtlcolor; `texture` before any mtlcolor
    b2 = list(lines); b2.insert(6, "sphere 0 0 0 1"); texts.append("\n".join(b2) + "\n")
    b2 = list(lines); b2.insert(6, f"texture {tex}"); b2.insert(9, "sphere 0 0 0 1"); texts.append("\n".join(b2) + "\n")
    outcomes = []
    for i, text in enumerate(texts):
        path = _write(tmp_path, text, f"d{i}.txt")
        want = _parse_with(path, 1, 1 << 30)
        outcomes.append(want[2] if want[0] == "err" else "ok")
        for threads, chunk in ((4, 64), (3, 257), (8, 4096)):
            got = _parse_with(path, threads, chunk)
            assert got == want, (i, threads, chunk, got[:3] if got[0] == "err" else "scene differs", want[:3] if want[0] == "err" else "")
    assert outcomes.count("ok") >= 4
    for needle in ("Invalid args for vertex", "'f' (face) object", "'sphere' object", "empty token", "Issue reading 'texture'",
                   "Invalid image dimensions", "Must define a 'mtlcolor'.", "Must define a 'mtlcolor' and 'texture'"):
        assert any(needle in o for o in outcomes), needle


def test_parser_missing_file(tmp_path):
    with pytest.raises(abi.RtError) as e:
        Scene.from_file(str(tmp_path / "absent.txt"))
    assert e.value.code == abi.RT_E_IO and "Issue reading input file" in e.value.message


def test_texture_loader(tmp_path):
    img = scenegen.make_texture("checker", 7, 5, 3)
    p = tmp_path / "t.ppm"
    p.write_text("P3\n# comment line\n7 5\n255\n" + "\n".join(" ".join(map(str, r.reshape(-1))) for r in img) + "\n")
    got = ppm.read_texture_p3(str(p))
    assert np.array_equal(got, img)
    # header tokens on one line, several spaces between values
    p.write_text("P3 7 5 255\n" + "  ".join(map(str, img.reshape(-1))) + "\n")
    assert np.array_equal(ppm.read_texture_p3(str(p)), img)
    for bad, needle in [("P6 7 5 255\n1 2 3\n", "P3"), ("P3 7 5 65535\n1 2 3\n", "255"), ("P3\n\n7 5\n255\n", "empty line"),
                        ("P3 2 2 255\n1 2 3\n", "too few"), ("P3 1 1 255\n1 2 300\n", "outside")]:
        p.write_text(bad)
        with pytest.raises(abi.RtError) as e:
            ppm.read_texture_p3(str(p))
        assert needle in e.value.message


def test_texture_loader_chunks_agree_with_one_pass(tmp_path):
    """The pixel part of a P3 file is parsed in parallel chunks (ppm_io.cpp): same pixels and same
    verdict as the single chunk, incl. the reference's order of failures — an empty line anywhere is
    fatal (it tokenises the whole file first), a bad token only among the first w*h*3."""
    rng = np.random.default_rng(11)
    img = rng.integers(0, 256, (23, 17, 3), dtype=np.uint8)
    rows = [" ".join(map(str, r.reshape(-1))) for r in img]
    body = "\n".join(rows) + "\n"
    files = {
        "plain": "P3\n17 23\n255\n" + body,
        "comments": "P3\n17 23\n255\n" + "\n".join(r + ("\n# note" if i % 5 == 0 else "") for i, r in enumerate(rows)) + "\n",
        "one_line_header": "P3 17 23 255 " + body,
        "crlf_values": "P3\n17 23\n255\n" + body.replace("\n", "\r\n"),
        "surplus_junk": "P3\n17 23\n255\n" + body + "these tokens are never converted 999 -1\n",
        "surplus_then_empty_line": "P3\n17 23\n255\n" + body + "1 2 3\n\n4 5 6\n",
        "bad_token_early": "P3\n17 23\n255\n" + body.replace(rows[3], rows[3].replace(" ", " x ", 1), 1),
        "bad_and_empty": "P3\n17 23\n255\n" + body.replace(rows[3], "300 1 2", 1) + "\n\n",
        "too_few": "P3\n17 23\n255\n" + "\n".join(rows[:20]) + "\n",
        "no_final_newline": "P3\n17 23\n255\n" + body.rstrip("\n"),
    }

    def load(path, threads, chunk):
        old = {k: os.environ.get(k) for k in ("RT_HOST_THREADS", "RT_HOST_CHUNK_BYTES")}
        os.environ["RT_HOST_THREADS"] = str(threads)
        os.environ["RT_HOST_CHUNK_BYTES"] = str(chunk)
        try:
            return ("ok", ppm.read_texture_p3(path).tobytes())
        except abi.RtError as e:
            return ("err", e.message)
        finally:
            for k, v in old.items():
                if v is None:
                    os.environ.pop(k, None)
                else:
                    os.environ[k] = v

    verdicts = {}
    for name, text in files.items():
        path = _write(tmp_path, text, name + ".ppm")
        want = load(path, 1, 1 << 30)
        verdicts[name] = want
        for threads, chunk in ((4, 50), (3, 333), (8, 4096)):
            assert load(path, threads, chunk) == want, (name, threads, chunk)
    for name in ("plain", "comments", "one_line_header", "crlf_values", "surplus_junk", "no_final_newline"):
        assert verdicts[name] == ("ok", img.tobytes()), name
    assert "empty line" in verdicts["surplus_then_empty_line"][1] and "empty line" in verdicts["bad_and_empty"][1]
    assert "stoi" in verdicts["bad_token_early"][1] and "too few" in verdicts["too_few"][1]


def test_ppm_writer_text_is_the_reference_dialect():
    rgb = np.array([[[0, 7, 255], [12, 100, 99]], [[1, 2, 3], [0, 0, 0]]], dtype=np.uint8)
    flags = np.array([[0, 0], [7, 8]], dtype=np.uint8)
    text = ppm.format_ppm(rgb, flags, bkg=(2.0, -0.5, 0.25))
    nan = "18446744071562067968"
    neg = str((1 << 64) - 127)                    # (size_t)(int)(-0.5 * 255)
    assert text == f"P3 \n2 2 \n255 \n0 7 255 \n12 100 99 \n{nan} {nan} {nan} \n510 {neg} 63 \n"


@pytest.mark.skipif(not os.path.exists(REF_BIN), reason="oracle/_ref/raytracer_ref not built")
@pytest.mark.parametrize("name", ["reflection_transparency_tests/Test1", "basic_geometry_tests/purple_pyramid"])
def test_ppm_writer_is_byte_identical_to_the_reference_program(tmp_path, name):
    """parser -> oracle frame -> our writer  ==  the reference program's own file."""
    scene, _ = common.golden_frame(name)
    text = common.scene_text(name)
    text = "\n".join(f"imsize {scene.width} {scene.height}" if l.startswith("imsize ") else l for l in text.split("\n"))
    sp = _write(tmp_path, text, "scene.txt")
    subprocess.run([REF_BIN, "scene.txt"], cwd=str(tmp_path), check=True, capture_output=True)
    ref_bytes = (tmp_path / "scene.ppm").read_bytes()
    parsed = Scene.from_file(sp)
    raw, ids, _ = O.oracle_render(parsed)
    rgb, flags = O.raw_to_rgb8(raw, ids)
    out = tmp_path / "ours.ppm"
    ppm.write_ppm(str(out), rgb.reshape(parsed.height, parsed.width, 3), flags.reshape(parsed.height, parsed.width), parsed.bkg)
    assert out.read_bytes() == ref_bytes
    assert ppm.output_path("dir.v2/scene.txt") == "dir.v2/scene.ppm"


def test_scene_text_round_trip(tmp_path):
    sc = scenegen.coverage_scene(32, 24)
    paths = []
    for i, t in enumerate(sc.textures):
        p = tmp_path / f"tex{i}.ppm"
        p.write_text(scenegen.p3_text(t))
        paths.append(str(p))
    back = Scene.from_file(_write(tmp_path, sc.to_text(paths)))
    raw_a, _, _ = O.oracle_render(sc)
    raw_b, _, _ = O.oracle_render(back)
    assert np.array_equal(raw_a, raw_b)


def test_render_without_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from simple_raytracer_b200.renderer import DeviceScene
    with pytest.raises(abi.RtError) as e:
        DeviceScene(scenegen.coverage_scene(16, 12))
    assert e.value.code == abi.RT_E_CUDA and "no CPU fallback" in e.value.message


def test_ctypes_mirror_has_the_sizes_the_c_compiler_sees(tmp_path):
    """sizeof / offsetof of the ABI structs as gcc lays them out from include/rt.h against the
    ctypes mirror (a mismatch in RtStats would let rt_render write past the Python object)."""
    import shutil
    import subprocess
    cc = shutil.which("gcc") or shutil.which("cc")
    if cc is None:
        pytest.skip("no C compiler")
    root = os.path.dirname(common.HERE)
    src = tmp_path / "sizes.c"
    src.write_text(
        '#include <stdio.h>\n#include <stddef.h>\n#include "rt.h"\n'
        "int main(void) {\n"
        '  printf("%zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(RtRenderParams), sizeof(RtStats), sizeof(RtSceneDesc),\n'
        "         sizeof(RtMaterial), sizeof(RtFace), offsetof(RtStats, spawn_ms), offsetof(RtRenderParams, work_sharing),\n"
        "         offsetof(RtStats, shadow_rays_walked));\n"
        "  return 0; }\n")
    exe = tmp_path / "sizes"
    subprocess.run([cc, "-std=c11", "-I", os.path.join(root, "include"), "-o", str(exe), str(src)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()]
    want = [C.sizeof(abi.RtRenderParams), C.sizeof(abi.RtStats), C.sizeof(abi.RtSceneDesc), C.sizeof(abi.RtMaterial),
            C.sizeof(abi.RtFace), abi.RtStats.spawn_ms.offset, abi.RtRenderParams.work_sharing.offset,
            abi.RtStats.shadow_rays_walked.offset]
    assert got == want
