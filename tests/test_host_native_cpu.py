"""C++ host runtime (csrc/host) and the TensorBoard writer built on it."""
import os
import struct
import zlib

import numpy as np
import torch

from bigdl_b200.utils import native


def test_crc_known_vectors():
    assert native.crc32c(b"123456789") == 0xE3069283          # RFC 3720 check value
    assert native.crc32c(b"") == 0
    assert native.crc32(b"123456789") == zlib.crc32(b"123456789")
    blob = os.urandom(100003)
    assert native.crc32(blob) == zlib.crc32(blob)
    # incremental == one-shot
    assert native.crc32c(blob[5000:], native.crc32c(blob[:5000])) == native.crc32c(blob)


def test_tfrecord_roundtrip_and_corruption():
    recs = [b"", b"a", os.urandom(1000)]
    buf = b"".join(native.tfrecord_frame(r) for r in recs)
    assert list(native.tfrecord_iter(buf)) == recs
    bad = bytearray(buf)
    bad[-10] ^= 0xFF
    try:
        list(native.tfrecord_iter(bytes(bad)))
        assert False
    except ValueError:
        pass


def test_bf16_codec_matches_bit_truncation():
    x = np.random.randn(200001).astype(np.float32)
    h = native.bf16_compress(x)
    assert np.array_equal(h, (x.view(np.uint32) >> 16).astype(np.uint16))
    y = native.bf16_decompress(h)
    assert np.array_equal(y.view(np.uint32), (x.view(np.uint32) >> 16) << 16)
    a, b = native.bf16_compress(x), native.bf16_compress(x[::-1].copy())
    want = native.bf16_compress(native.bf16_decompress(a) + native.bf16_decompress(b))
    assert np.array_equal(native.bf16_add_(a, b), want)


def test_native_rng_matches_python_stream():
    from bigdl_b200.utils.random_generator import RandomGenerator

    r = native.NativeRNG(42)
    py = RandomGenerator().setSeed(42)
    a = [r.uniform(0, 1) for _ in range(5)]
    b = [py.uniform(0, 1) for _ in range(5)]
    assert np.allclose(a, b)
    perm = native.NativeRNG(7).permutation(1000)
    assert sorted(perm.tolist()) == list(range(1000))


def test_batch_images_matches_numpy():
    rng = np.random.RandomState(0)
    imgs = [rng.randint(0, 256, size=(40 + i, 50 + i, 3), dtype=np.uint8) for i in range(9)]
    mean, std = [0.4, 0.5, 0.6], [0.2, 0.25, 0.3]
    out = native.batch_images(imgs, 32, 32, mean, std, scale=255.0)
    for i, im in enumerate(imgs):
        oy, ox = (im.shape[0] - 32) // 2, (im.shape[1] - 32) // 2
        ref = (im[oy:oy + 32, ox:ox + 32].astype(np.float32) / 255.0 - mean) / std
        assert np.allclose(out[i], ref.transpose(2, 0, 1), atol=1e-5)
    # bf16 / NHWC straight into a pinned-style torch buffer
    buf = torch.empty(9, 32, 32, 3, dtype=torch.bfloat16)
    native.batch_images(imgs, 32, 32, mean, std, scale=255.0, nhwc=True, bf16=True, out=buf)
    assert torch.allclose(buf.float(), torch.from_numpy(out).permute(0, 2, 3, 1), atol=2e-2)
    # seeded random crop + flip is deterministic
    seeds = np.arange(9, dtype=np.uint64)
    a = native.batch_images(imgs, 24, 24, mean, std, random_crop=True, random_flip=True, seeds=seeds)
    b = native.batch_images(imgs, 24, 24, mean, std, random_crop=True, random_flip=True, seeds=seeds)
    assert np.array_equal(a, b)


def test_int8_reference_gemm():
    rng = np.random.RandomState(1)
    a, b = rng.randn(17, 64).astype(np.float32), rng.randn(9, 64).astype(np.float32)
    bias = rng.randn(9).astype(np.float32)
    aq, aam = native.quantize_rows_int8(a)
    bq, bam = native.quantize_rows_int8(b)
    assert np.array_equal(aq, np.clip(np.rint(a / aam[:, None] * 127), -127, 127).astype(np.int8))
    c = native.int8_gemm(aq, aam, bq, bam, bias)
    ref = (aq.astype(np.int32) @ bq.astype(np.int32).T) * (aam[:, None] / 127) * (bam[None, :] / 127) + bias
    assert np.allclose(c, ref, rtol=1e-5, atol=1e-5)
    assert np.abs(c - (a @ b.T + bias)).max() < 0.3


def test_train_summary_roundtrip(tmp_path):
    from bigdl_b200.optim.trigger import Trigger
    from bigdl_b200.visualization import TrainSummary, ValidationSummary

    ts = TrainSummary(str(tmp_path), "app")
    for i in range(10):
        ts.addScalar("Loss", 1.0 / (i + 1), i)
    ts.addHistogram("fc/weight", torch.randn(1000), 3)
    got = ts.readScalar("Loss")
    assert [g[0] for g in got] == list(range(10))
    assert np.allclose([g[1] for g in got], [1.0 / (i + 1) for i in range(10)])
    ts.setSummaryTrigger("Parameters", Trigger.severalIteration(2))
    assert ts.has_trigger("Parameters") and not ts.has_trigger("gradientNorm2")
    try:
        ts.setSummaryTrigger("Bogus", Trigger.everyEpoch())
        assert False
    except ValueError:
        pass
    ts.close()
    vs = ValidationSummary(str(tmp_path), "app")
    vs.addScalar("Top1Accuracy", 0.5, 1)
    assert vs.readScalar("Top1Accuracy")[0][1] == 0.5
    vs.close()
    # the event file starts with a file_version record, framed as TFRecord
    fn = [f for f in os.listdir(os.path.join(str(tmp_path), "app", "train"))][0]
    data = open(os.path.join(str(tmp_path), "app", "train", fn), "rb").read()
    ln = struct.unpack("<Q", data[:8])[0]
    assert b"brain.Event:2" in data[12:12 + ln]


def test_optimizer_writes_summaries(tmp_path):
    from bigdl_b200.dataset.base import DataSet, Sample
    from bigdl_b200.nn import layer as L
    from bigdl_b200.nn.criterion import MSECriterion
    from bigdl_b200.optim.optim_method import SGD
    from bigdl_b200.optim.optimizer import Optimizer
    from bigdl_b200.optim.trigger import Trigger
    from bigdl_b200.visualization import TrainSummary

    torch.manual_seed(0)
    samples = [Sample(np.random.randn(4).astype(np.float32), np.random.randn(1).astype(np.float32)) for _ in range(32)]
    model = L.Sequential().add(L.Linear(4, 1))
    opt = Optimizer(model=model, training_set=DataSet.array(samples), criterion=MSECriterion(), optim_method=SGD(0.01),
                    end_trigger=Trigger.maxIteration(6), batch_size=8)
    ts = TrainSummary(str(tmp_path), "run")
    ts.setSummaryTrigger("Parameters", Trigger.severalIteration(3))
    opt.setTrainSummary(ts)
    opt.optimize()
    assert len(ts.readScalar("Loss")) == 6 and len(ts.readScalar("LearningRate")) == 6
    ts.close()


def test_affinity_and_core_counts():
    from bigdl_b200.utils import native
    cur = native.get_affinity()
    assert cur and all(isinstance(c, int) for c in cur)
    assert 1 <= native.physical_cores() <= native.logical_cores()
    native.set_affinity(cur[:1])
    try:
        assert native.get_affinity() == cur[:1]
    finally:
        native.set_affinity(cur)
    assert native.get_affinity() == cur
