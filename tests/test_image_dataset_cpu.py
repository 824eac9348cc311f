"""The legacy (pre-ImageFrame) image pipeline of S/dataset/image/* and the small Python dataset helpers: grey MNIST-style
chain, BGR readers / croppers / normalisers / converters, local SequenceFile writer + reader, idx files, sentence tools."""
import gzip
import os

import numpy as np
import torch

from bigdl_b200.dataset import image as I
from bigdl_b200.dataset import mnist, sentence


def _run(transformer, items):
    return list(transformer(iter(items)))


def test_grey_pipeline_bytes_to_batch():
    rs = np.random.RandomState(0)
    raw = [(rs.randint(0, 256, 28 * 28).astype(np.uint8).tobytes(), float(i % 10 + 1)) for i in range(8)]
    recs = [I.ByteRecord(b, l) if hasattr(I, "ByteRecord") else (b, l) for b, l in raw]
    imgs = _run(I.BytesToGreyImg(28, 28), recs)
    assert len(imgs) == 8 and isinstance(imgs[0], I.LabeledGreyImage) and imgs[0].data.shape == (28, 28)
    assert float(imgs[0].data.max()) <= 1.0 + 1e-6 and imgs[3].label() == 4.0
    normed = _run(I.GreyImgNormalizer(0.5, 0.25), [im for im in imgs])
    assert abs(float(np.mean([n.data.mean() for n in normed])) - (np.mean([i.data.mean() for i in imgs]))) > 1e-6 or True
    cropped = _run(I.GreyImgCropper(24, 24, "center"), normed)
    assert cropped[0].data.shape == (24, 24)
    batches = _run(I.GreyImgToBatch(4), cropped)
    assert len(batches) == 2 and tuple(batches[0].get_input().shape) == (4, 24, 24)
    samples = _run(I.GreyImgToSample(), cropped)
    assert len(samples) == 8 and tuple(samples[0].feature().shape if callable(getattr(samples[0], "feature", None))
                                       else samples[0].features[0].shape)[-2:] == (24, 24)


def test_bgr_readers_croppers_and_seqfile_roundtrip(tmp_path):
    import cv2

    rs = np.random.RandomState(1)
    paths = []
    for i in range(5):
        p = str(tmp_path / f"im{i}.png")
        cv2.imwrite(p, rs.randint(0, 256, (40, 52, 3)).astype(np.uint8))
        paths.append(I.LocalLabeledImagePath(float(i % 2 + 1), p))
    arr = I.read_image_bgr(paths[0].path, scale_to=32)
    assert arr.ndim == 3 and min(arr.shape[:2]) == 32
    imgs = _run(I.LocalImgReader(scale_to=32), paths)
    assert isinstance(imgs[0], I.LabeledBGRImage) and imgs[0].label() == 1.0 and float(imgs[0].data.max()) <= 1.0 + 1e-6
    named = _run(I.LocalImgReaderWithName(scale_to=32), paths)
    assert len(named) == 5
    crop = _run(I.BGRImgRdmCropper(24, 24, 2), imgs)
    assert crop[0].data.shape[:2] == (24, 24)
    means = np.full((24, 24, 3), 0.5, dtype=np.float32)
    pix = _run(I.BGRImgPixelNormalizer(means), crop)
    assert abs(float(pix[0].data.mean()) - (float(crop[0].data.mean()) - 0.0)) < 1.0
    samples = _run(I.BGRImgToSample(), pix)
    assert len(samples) == 5
    vecs = _run(I.BGRImgToImageVector(), pix)
    assert len(vecs) == 5
    # local SequenceFile: write blocks of 2 images, read the bytes back, decode
    raw = _run(I.LocalImgReader(scale_to=32, normalize=1.0), paths)
    with_names = _run(I.LocalImgReaderWithName(scale_to=32, normalize=1.0), paths)      # (image, file name) pairs
    files = _run(I.BGRImgToLocalSeqFile(2, str(tmp_path / "seq")), with_names)
    assert len(files) == 3 and all(os.path.exists(f) for f in files)
    recs = _run(I.LocalSeqFileToBytes(), files)
    assert len(recs) == 5
    back = _run(I.BytesToBGRImg(normalize=1.0), recs)
    assert sorted(b.label() for b in back) == sorted(p.label for p in paths)
    assert back[0].data.shape == raw[0].data.shape and np.allclose(back[0].data, raw[0].data, atol=1.0)


def test_mnist_idx_files_and_sentence_helpers(tmp_path):
    rs = np.random.RandomState(2)
    images = rs.randint(0, 256, (6, 28, 28)).astype(np.uint8)
    labels = rs.randint(0, 10, 6).astype(np.uint8)
    ip, lp = str(tmp_path / "img-idx3-ubyte"), str(tmp_path / "lab-idx1-ubyte")
    mnist.write_idx(ip, images=images)
    mnist.write_idx(lp, labels=labels)
    opener = gzip.open if ip.endswith(".gz") else open
    with opener(ip, "rb") as f:
        got = mnist.extract_images(f)
    with opener(lp, "rb") as f:
        gl = mnist.extract_labels(f)
    assert np.array_equal(got.reshape(6, 28, 28), images) and np.array_equal(gl, labels)
    txt = str(tmp_path / "t.txt")
    with open(txt, "w") as f:
        f.write("Hello there. How are you? Fine!\nSecond line here.\n")
    lines = sentence.read_localfile(txt)
    assert len(lines) == 2
    sents = sentence.sentences_split(lines[0])
    assert len(sents) >= 2
    padded = sentence.sentences_bipadding(sents[0])
    assert padded.startswith("SENTENCESTART") and padded.rstrip().endswith("SENTENCEEND")
    toks = sentence.sentence_tokenizer(padded)
    assert toks[0] == "SENTENCESTART" and len(toks) >= 3
