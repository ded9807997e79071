"""the Middlebury on-disk format (reference stereomatch/data.py:14-93) and
the quality metrics of the evaluation step (reference workflows/evaluation/workflow.py:74-105),
on a synthetic scene written to a temp dir in that layout.  No Middlebury data is on this box."""
import numpy as np
import pytest
import torch

from oracle import synth


def write_scene(root, name, H, W, D, seed, rgb=True):
    from PIL import Image
    from stereomatch.data import write_pfm
    left, right, gt = synth.random_dot_pair(H, W, D, seed=seed, scale="raw")
    d = root / name
    d.mkdir(parents=True)
    for fname, img in (("im0.png", left), ("im1.png", right)):
        u8 = img.astype(np.uint8)
        Image.fromarray(np.stack([u8] * 3, axis=2) if rgb else u8).save(d / fname)
    gtf = gt.astype(np.float32)
    gtf[:, : D // 2] = np.inf                      # unknown disparities, as Middlebury marks them
    write_pfm(d / "disp0.pfm", gtf)
    write_pfm(d / "disp1.pfm", gtf)
    (d / "calib.txt").write_text(
        f"cam0=[1733.74 0 792.27; 0 1733.74 541.89; 0 0 1]\ncam1=[1733.74 0 792.27; 0 1733.74 541.89; 0 0 1]\n"
        f"doffs=0\nbaseline=536.62\nwidth={W}\nheight={H}\nndisp={D - 3}\nvmin=5\nvmax={D}\n")
    return left, right, gtf


def test_pfm_round_trip_and_byte_order(tmp_path):
    from stereomatch.data import read_pfm, write_pfm
    rng = np.random.default_rng(0)
    a = rng.random((7, 13)).astype(np.float32) * 100
    a[2, 3] = np.inf
    write_pfm(tmp_path / "a.pfm", a)
    assert np.array_equal(read_pfm(tmp_path / "a.pfm"), a)
    raw = (tmp_path / "a.pfm").read_bytes()
    assert raw.startswith(b"Pf\n13 7\n-1.0\n")
    # rows are stored bottom first
    assert np.frombuffer(raw[-13 * 4:], "<f4").tolist() == a[0].tolist()
    # big-endian file (positive scale), colour
    c = rng.random((4, 5, 3)).astype(np.float32)
    with open(tmp_path / "c.pfm", "wb") as f:
        f.write(b"PF\n5 4\n1.0\n")
        f.write(c[::-1].astype(">f4").tobytes())
    assert np.array_equal(read_pfm(tmp_path / "c.pfm"), c)
    write_pfm(tmp_path / "c2.pfm", c)
    assert np.array_equal(read_pfm(tmp_path / "c2.pfm"), c)
    (tmp_path / "bad.pfm").write_bytes(b"P5\n1 1\n255\n\0")
    with pytest.raises(ValueError):
        read_pfm(tmp_path / "bad.pfm")
    (tmp_path / "short.pfm").write_bytes(b"Pf\n4 4\n-1\n" + b"\0" * 12)
    with pytest.raises(ValueError):
        read_pfm(tmp_path / "short.pfm")


def test_pfm_matches_opencv_reader(tmp_path):
    """The reference reads ground truth with cv2.imread(IMREAD_UNCHANGED) (data.py:86)."""
    cv2 = pytest.importorskip("cv2")
    from stereomatch.data import read_pfm, write_pfm
    a = (np.arange(6 * 9, dtype=np.float32).reshape(6, 9) * 0.37)
    a[1, 1] = np.inf
    write_pfm(tmp_path / "g.pfm", a)
    assert np.array_equal(cv2.imread(str(tmp_path / "g.pfm"), cv2.IMREAD_UNCHANGED), read_pfm(tmp_path / "g.pfm"))


def test_middlebury_dataset_layout(tmp_path):
    from stereomatch.data import MiddleburyDataset, _parse_middlebury_calib
    root = tmp_path / "data"
    scenes = {}
    for name, seed in (("scene10", 3), ("scene2", 4), ("artroom1", 5)):
        scenes[name] = write_scene(root, name, 24, 40, 16, seed, rgb=name != "scene2")
    (root / "README.txt").write_text("not a scene")
    ds = MiddleburyDataset(str(root))
    assert len(ds) == 3
    assert [p[0].parent.name for p in ds.images] == ["artroom1", "scene2", "scene10"]   # natural order
    assert _parse_middlebury_calib(root / "scene2" / "calib.txt") == dict(width=40, height=24, ndisp=13)
    item = ds[2]
    left, right, gt = scenes["scene10"]
    assert item["stereo_name"] == "scene10" and item["max_disparity"] == 13
    assert item["left"].dtype == torch.uint8 and tuple(item["left"].shape) == (24, 40, 3)
    assert np.array_equal(item["left"][..., 0].numpy(), left.astype(np.uint8))
    assert np.array_equal(item["right"][..., 1].numpy(), right.astype(np.uint8))
    assert item["gt_disparity"].dtype == torch.float32 and np.array_equal(item["gt_disparity"].numpy(), gt)
    assert tuple(ds[1]["left"].shape) == (24, 40)                                        # grayscale file
    # as in the reference (data.py:47-51) max_size cuts the directory listing BEFORE non-directories
    # are skipped: the first two entries are artroom1 and README.txt
    assert len(MiddleburyDataset(str(root), max_size=2)) == 1
    assert len(MiddleburyDataset(str(root), max_size=3)) == 2
    with pytest.raises(RuntimeError, match="must be a directory"):
        MiddleburyDataset(str(root / "README.txt"))


def test_metrics_definitions():
    from stereomatch import metrics
    gt = torch.tensor([[1.0, 2.0, float("inf")], [4.0, 10.0, 6.0]])
    pred = torch.tensor([[1, 5, 100], [4, 7, 6]], dtype=torch.int32)
    # valid diffs: 0 3 0 3 0  ->  rmse sqrt(18/5), avgerr 6/5, bad(2.0) 2/5
    r = metrics.evaluate(gt, pred)
    assert r["valid_pixels"] == 5
    assert r["rmse"] == pytest.approx((18 / 5) ** 0.5) and r["avgerr"] == pytest.approx(1.2)
    assert r["bad_pix_error"] == pytest.approx(40.0)
    assert metrics.calc_bad_pix_error(gt, pred, threshold=3.0) == 0.0           # strictly greater
    assert metrics.calc_rmse(gt, pred) == r["rmse"] and metrics.calc_avgerr(gt, pred) == r["avgerr"]
    mask = torch.tensor([[True, False, True], [True, True, True]])
    assert metrics.evaluate(gt, pred, mask=mask)["valid_pixels"] == 4
    allinf = torch.full((2, 2), float("inf"))
    assert metrics.evaluate(allinf, torch.zeros(2, 2))["rmse"] == 0.0           # nothing valid: no NaN
    with pytest.raises(ValueError):
        metrics.evaluate(gt, pred[:1])
    # against a numpy restatement on random data
    rng = np.random.default_rng(1)
    g = rng.random((50, 60)).astype(np.float32) * 64
    g[rng.random((50, 60)) < 0.2] = np.inf
    p = rng.integers(0, 64, (50, 60)).astype(np.int32)
    v = np.isfinite(g)
    d = np.abs(p[v].astype(np.float64) - g[v])
    r = metrics.evaluate(torch.from_numpy(g), torch.from_numpy(p), threshold=1.5)
    assert r["rmse"] == pytest.approx(np.sqrt((d * d).mean())) and r["avgerr"] == pytest.approx(d.mean())
    assert r["bad_pix_error"] == pytest.approx(100.0 * (d > 1.5).mean())


def test_grayscale_is_the_torchvision_formula():
    from stereomatch.evaluation import to_grayscale_f32
    rgb = torch.tensor([[[255, 0, 0], [0, 255, 0], [0, 0, 255], [10, 20, 30]]], dtype=torch.uint8)
    want = torch.tensor([[int(0.2989 * 255), int(0.587 * 255), int(0.114 * 255),
                          int(0.2989 * 10 + 0.587 * 20 + 0.114 * 30)]]).float()
    assert torch.equal(to_grayscale_f32(rgb), want)
    assert torch.equal(to_grayscale_f32(rgb[..., 0]), rgb[..., 0].float())


@pytest.mark.gpu
def test_evaluate_dataset_scores_the_pipeline_on_device(tmp_path):
    """Predict + score a two-scene synthetic dataset; the random-dot ground truth is recovered
    (SURVEY.md 8(d) anchors: SSD+WTA / SSD+SGM+WTA > 99 % exact where a match exists)."""
    import stereomatch
    from stereomatch.cli_common import create_pipeline
    from stereomatch.data import MiddleburyDataset
    from stereomatch.evaluation import evaluate_dataset
    root = tmp_path / "data"
    for name, seed in (("a1", 11), ("a2", 12)):
        write_scene(root, name, 96, 160, 32, seed)
    ds = MiddleburyDataset(str(root))
    for cost, red, agg in (("ssd", "wta", None), ("ssd", "wta", "sgm"), ("ssd-texture", "wta", None),
                           ("birchfield", "wta", "sgm"), ("ssd", "dyn", None)):
        table = evaluate_dataset(ds, create_pipeline(cost, red, agg, 32))
        assert len(table["scenes"]) == 2 and table["scenes"][0]["stereo_name"] == "a1"
        assert table["scenes"][0]["valid_pixels"] == 96 * (160 - 16)
        print(f"{cost}-{red}-{agg}: rmse {table['rmse']:.3f} avgerr {table['avgerr']:.3f} "
              f"bad2.0 {table['bad_pix_error']:.2f} %")
        if red == "wta":
            assert table["bad_pix_error"] < 5.0 and table["avgerr"] < 1.0
    # the metrics ran where the prediction lives
    pipe = create_pipeline("ssd", "wta", "sgm", 32)
    item = ds[0]
    l = item["left"][..., 0].float().cuda()
    r = item["right"][..., 0].float().cuda()
    disp = pipe.estimate(l, r)
    res = stereomatch.metrics.evaluate(item["gt_disparity"].cuda(), disp)
    assert res["valid_pixels"] == 96 * (160 - 16) and res["bad_pix_error"] < 5.0
