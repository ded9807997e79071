"""The image command line: string tables, file I/O, and -- on a GPU -- that
the CLI writes exactly what Pipeline.estimate returns.  Interface follows reference
stereomatch/cli_common.py:9-52 and cli_image.py:34-62."""
import numpy as np
import pytest
import torch

from oracle import synth


@pytest.fixture(scope="module")
def sm():
    import stereomatch
    return stereomatch


def test_string_tables_match_the_reference(sm):
    from stereomatch import cli_common as cc
    assert list(cc.COST_METHODS) == ["ssd", "ssd-texture", "birchfield"]      # cli_common.py:9-13
    assert list(cc.AGGREGATION_METHODS) == ["sgm"]                            # :15-17
    assert list(cc.DISPARITY_METHODS) == ["wta", "dyn"]                       # :19-22


def test_create_pipeline_builds_the_named_stages(sm):
    from stereomatch.cli_common import create_pipeline
    p = create_pipeline("ssd", "wta", "sgm", max_disparity=48)
    assert type(p.cost) is sm.cost.SSD and p.cost.max_disparity == 48
    assert type(p.aggregation) is sm.aggregation.Semiglobal
    assert type(p.disparity_reduce) is sm.disparity_reduce.WinnerTakesAll
    q = create_pipeline("birchfield", "dyn", None)
    assert type(q.cost) is sm.cost.Birchfield and q.cost.max_disparity == 32   # default, :26
    assert q.aggregation is None and type(q.disparity_reduce) is sm.disparity_reduce.DynamicProgramming
    assert create_pipeline("ssd", "wta", "no-such-aggregation").aggregation is None   # .get(), :46
    with pytest.raises(KeyError):
        create_pipeline("census", "wta", None)
    t = create_pipeline("ssd-texture", "wta", None)     # cli_common.py:11; Pipeline wraps it, pipeline.py:54-55
    assert type(t.cost).__name__ == "_TexCostFunctionWrapper"
    assert type(t.cost.cost_function) is sm.cost.SSDTexture and t.cost.cost_function.max_disparity == 32


@pytest.mark.parametrize("ext", ["png", "pgm"])
def test_u16_files_round_trip(sm, tmp_path, ext):
    from stereomatch.cli_image import read_u16, write_u16
    d = (np.arange(7 * 13, dtype=np.int32).reshape(7, 13) * 311) % 60000   # needs 16 bits
    path = str(tmp_path / f"d.{ext}")
    write_u16(path, d)
    back = read_u16(path)
    assert back.dtype == np.uint16 and np.array_equal(back, d.astype(np.uint16))


def test_parser_defaults_and_options(sm):
    from stereomatch.cli_image import build_parser
    a = build_parser().parse_args(["l.png", "r.png", "64", "out.png"])
    assert (a.cost_method, a.aggregation_method, a.disparity_method, a.max_disparity) == ("ssd", None, "wta", 64)
    b = build_parser().parse_args(["l.png", "r.png", "128", "o.pgm", "-cm", "birchfield", "-am", "sgm",
                                   "-dm", "dyn", "-c", "-sd", "-fig"])
    assert (b.cost_method, b.aggregation_method, b.disparity_method) == ("birchfield", "sgm", "dyn")


def test_gray_conversion_is_pil_luma(sm, tmp_path):
    from PIL import Image
    from stereomatch.cli_image import read_gray_f32
    rgb = (np.random.default_rng(3).integers(0, 256, (9, 11, 3))).astype(np.uint8)
    Image.fromarray(rgb).save(tmp_path / "c.png")
    got = read_gray_f32(str(tmp_path / "c.png"))
    want = np.array(Image.fromarray(rgb).convert("L")).astype(np.float32)
    assert got.dtype == torch.float32 and np.array_equal(got.numpy(), want)


@pytest.mark.gpu
@pytest.mark.parametrize("opts", [["-cm", "ssd", "-am", "sgm", "-dm", "wta"], ["-cm", "birchfield", "-dm", "dyn"],
                                  ["-cm", "ssd"], ["-cm", "ssd-texture", "-am", "sgm"]])
def test_cli_writes_what_the_pipeline_returns(sm, tmp_path, opts):
    from PIL import Image
    from stereomatch import cli_image
    from stereomatch.cli_common import create_pipeline
    H, W, D = 40, 72, 32
    left, right, _ = synth.random_dot_pair(H, W, D, seed=21, scale="raw")
    for name, img in (("l.png", left), ("r.png", right)):
        Image.fromarray(img.astype(np.uint8)).save(tmp_path / name)
    out = str(tmp_path / "disp.pgm")
    assert cli_image.main([str(tmp_path / "l.png"), str(tmp_path / "r.png"), str(D), out, "-c"] + opts) == 0
    ns = cli_image.build_parser().parse_args(["a", "b", str(D), "c"] + opts)
    pipe = create_pipeline(ns.cost_method, ns.disparity_method, ns.aggregation_method, D)
    want = pipe.estimate(torch.from_numpy(left).cuda(), torch.from_numpy(right).cuda()).cpu().numpy()
    got = cli_image.read_u16(out)
    assert np.array_equal(got, want.astype(np.uint16))
    # ... and what it writes is the ORACLE's chain on the same files (<= 0.1 % of pixels, north_star)
    from oracle import oracle as O
    vol = O.ssd(left, right, D, 7) if ns.cost_method.startswith("ssd") else O.birchfield(left, right, D)
    if ns.aggregation_method == "sgm":
        vol = O.sgm(vol, left, 0.1, 0.2)
    owant = O.wta(vol) if ns.disparity_method == "wta" else O.dp(vol, "cuda")
    assert float((got.astype(np.int32) != owant).mean()) <= 1e-3
