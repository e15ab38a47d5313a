"""JPEG decode on the device (nvJPEG behind `evr_jpeg_decode_batch`) against PIL / libjpeg-turbo,
the decoder the reference uses (`Image.open(p).convert('RGB')`, Backend/services/embedding_service.py:475).

The two decoders are both valid JPEG decoders but not bit-identical (IDCT arithmetic, chroma up-sampling filter), so
the parity statement is in two parts, both written here:
  pixels      mean |nvJPEG - PIL| <= 2.5 grey levels and 99.9 % of the samples within 24 (real key-frames: mean 0.2); the
              maximum is only bounded loosely (<= 128): at sharp chroma edges of 4:2:0 streams libjpeg-turbo's "fancy"
              triangle up-sampling and nvJPEG's differ by tens of levels on isolated pixels
  embeddings  cosine >= 0.999 per frame between the rows ingested with `gpu_jpeg=True` and with PIL (north_star's
              tolerance), on the reference's own stored 1280x720 key-frames (tests/golden/ref_frames) and synthetic ones.
PIL remains the exact mode (default); `gpu_jpeg` is opt-in.
"""
import importlib
import json
import os
import time

import numpy as np
import pytest
import torch

from conftest import PKG

pytestmark = pytest.mark.gpu
FRAMES = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_frames")


def _real_frames():
    return sorted(os.path.join(FRAMES, f) for f in os.listdir(FRAMES) if f.endswith(".jpg"))


def _close(diff):
    return diff.mean() <= 2.5 and np.percentile(diff, 99.9) <= 24 and diff.max() <= 128


def _pil(path):
    from PIL import Image
    return np.array(Image.open(path).convert("RGB"))


@pytest.fixture(scope="module")
def decoder(pkg, cuda):
    dec = pkg.preprocess.JpegDecoder(cuda)
    print("nvJPEG backend:", dec.backend)
    return dec


def test_reference_keyframes_decode_close_to_pil(decoder):
    paths = _real_frames()
    assert len(paths) == 8
    out = decoder.decode_files(paths)
    assert out.shape == (8, 720, 1280, 3) and out.dtype == torch.uint8
    got = out.cpu().numpy().astype(np.int16)
    for i, p in enumerate(paths):
        diff = np.abs(got[i] - _pil(p).astype(np.int16))
        assert _close(diff), (p, int(diff.max()), float(diff.mean()), float(np.percentile(diff, 99.9)))
    print("real key-frames: max |diff|", int(np.abs(got[0] - _pil(paths[0])).max()))


@pytest.mark.parametrize("subsampling,quality", [(0, 95), (2, 75), (1, 85)])
def test_synthetic_jpegs(decoder, tmp_path, subsampling, quality):
    from PIL import Image
    rng = np.random.default_rng(subsampling * 10 + quality)
    paths = []
    for i in range(5):
        low = rng.integers(0, 255, (9, 16, 3), dtype=np.uint8)
        img = Image.fromarray(low).resize((320, 180), Image.BICUBIC)
        p = str(tmp_path / f"{i}.jpg")
        img.save(p, quality=quality, subsampling=subsampling)
        paths.append(p)
    got = decoder.decode_files(paths).cpu().numpy().astype(np.int16)
    for i, p in enumerate(paths):
        diff = np.abs(got[i] - _pil(p).astype(np.int16))
        assert _close(diff), (int(diff.max()), float(diff.mean()), float(np.percentile(diff, 99.9)))
    assert decoder.size(open(paths[0], "rb").read()) == (320, 180)


def test_mixed_sizes_and_garbage_are_refused(pkg, decoder, tmp_path):
    from PIL import Image
    a, b = str(tmp_path / "a.jpg"), str(tmp_path / "b.jpg")
    Image.new("RGB", (64, 48), (10, 200, 30)).save(a)
    Image.new("RGB", (80, 48), (10, 200, 30)).save(b)
    with pytest.raises(pkg._lib.EvrError, match="batch is"):
        decoder.decode_files([a, b])
    with pytest.raises(pkg._lib.EvrError):
        decoder.decode([b"not a jpeg at all"], size=(64, 48))
    ok = decoder.decode_files([a])          # the handle survives a failed batch
    assert ok.shape == (1, 48, 64, 3) and abs(int(ok[0, 0, 0, 1]) - 200) <= 2


def test_ingest_with_device_jpeg_matches_pil_ingest(tmp_path, cuda):
    """`EmbeddingService(..., gpu_jpeg=True)` over a folder holding the reference's real key-frames, a PNG (falls back to
    PIL) and a damaged file (zero row): rows agree with the PIL ingest to cosine >= 0.999; same layout, same order."""
    import shutil
    from PIL import Image
    services = importlib.import_module(PKG + ".services")
    clip = importlib.import_module(PKG + ".clip")
    folder = tmp_path / "frames" / "vid"
    folder.mkdir(parents=True)
    names = []
    for i, p in enumerate(_real_frames()):
        shutil.copyfile(p, folder / f"{100 + i}.jpg")
        names.append(f"{100 + i}.jpg")
    Image.fromarray(np.random.default_rng(1).integers(0, 255, (720, 1280, 3), dtype=np.uint8)).save(folder / "300.png")
    (folder / "200.jpg").write_bytes(b"\xff\xd8 broken")
    names += ["300.png", "200.jpg"]
    meta, emb = tmp_path / "metadata", tmp_path / "embedding"
    meta.mkdir()
    emb.mkdir()
    items = [{"filepath": str(folder / n), "frameidx": int(n.split(".")[0])} for n in sorted(names)]
    (meta / "vid_metadata.json").write_text(json.dumps(items))
    (meta / "video_mapping.json").write_text(json.dumps({"vid": {
        "metadata_file": str(meta / "vid_metadata.json"), "embeddings_file": str(emb / "vid_embeddings.npy"),
        "video_path": "", "frames_dir": str(folder)}}))
    sd = clip.random_visual_state_dict("ViT-B/32", seed=1)
    rows = {}
    for mode in (False, True):
        ps = services.PathService(str(tmp_path / "frames"), str(meta), str(emb))
        cs = services.CacheService()
        svc = services.EmbeddingService(cs, ps, services.DataService(ps, cs), "cuda",
                                        load_kwargs=dict(state_dict=sd, max_batch=16), gpu_jpeg=mode)
        t0 = time.perf_counter()
        path = svc.extract_and_save_embeddings_from_folder(str(folder), None, "vid")
        rows[mode] = np.load(path).astype(np.float32)
        print(f"gpu_jpeg={mode}: {len(names)} frames in {time.perf_counter() - t0:.3f} s")
    a, b = rows[False], rows[True]
    order = sorted(names)
    assert a.shape == b.shape == (10, 512)
    broken = order.index("200.jpg")
    assert np.all(a[broken] == 0) and np.all(b[broken] == 0)
    png = order.index("300.png")
    assert np.array_equal(a[png], b[png])                      # the PNG went through PIL in both runs
    live = [i for i in range(10) if i != broken]
    cos = (a[live] * b[live]).sum(-1) / (np.linalg.norm(a[live], axis=-1) * np.linalg.norm(b[live], axis=-1))
    print("cosine nvJPEG-vs-PIL ingest:", cos.round(6))
    assert cos.min() >= 0.999, cos
