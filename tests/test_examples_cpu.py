"""The bundled examples and CLI tools run end to end on CPU with synthetic data (the reference keeps its examples under
integration tests: spark/dl/src/test/integration-test.robot)."""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load(name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, "examples", name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def test_lenet_local_example_learns():
    assert _load("lenet_local").main(["-e", "3", "-f", "/nonexistent"]) > 0.9


def test_load_model_example_round_trips_every_format():
    assert _load("load_model").main([]) < 1e-5


def test_int8_inference_example():
    rel, agree = _load("int8_inference").main(["--iters", "2", "-o", "/tmp/_q_example.bigdl"])
    assert rel < 0.05 and agree > 0.8


def test_ml_pipeline_example():
    m = _load("ml_pipeline")
    assert m.logistic_regression() > 0.9 and m.multi_label_regression() < 0.05


def test_cli_tools(tmp_path, capsys):
    import json

    from bigdl_b200.cli import convert_model, imagenet_seqfile_generator, local_optimizer_perf

    local_optimizer_perf.main(["-m", "lenet5", "-b", "16", "-i", "12", "--device", "cpu"])
    line = [l for l in capsys.readouterr().out.splitlines() if l.startswith("{")][-1]
    assert json.loads(line)["records_per_second"] > 0
    # ImageNet folder → sequence files → dataset
    import cv2
    import numpy as np

    for split in ("train", "val"):
        for cls in ("n01", "n02"):
            os.makedirs(tmp_path / "imgs" / split / cls)
            for i in range(3):
                cv2.imwrite(str(tmp_path / "imgs" / split / cls / f"{i}.jpg"),
                            np.random.randint(0, 255, (40, 60, 3), dtype=np.uint8))
    imagenet_seqfile_generator.main(["-f", str(tmp_path / "imgs"), "-o", str(tmp_path / "seq"), "-p", "2", "-b", "4",
                                     "--scaleSize", "32", "-n"])
    from bigdl_b200.dataset.seqfile import seq_file_dataset

    ds = seq_file_dataset(str(tmp_path / "seq" / "train"), 1000, distributed=False)
    assert ds.size() == 6


def test_tree_lstm_sentiment_example_learns():
    assert _load("tree_lstm_sentiment").main(["-e", "3"]) > 0.7


def test_keras_imdb_lstm_example_learns():
    assert _load("keras_imdb_lstm").main(["-e", "3"]) > 0.85


def test_image_validation_example_runs_the_imageframe_path():
    assert _load("image_validation").main(["-n", "90", "-e", "25"]) > 0.9


def test_udf_predictor_example():
    assert _load("udf_predictor").main(["--maxEpoch", "5"]) > 0.8


def test_dlframes_image_example():
    acc, same = _load("dlframes_image").main([])
    assert acc >= 0.8 and same == 1.0


def test_tensorflow_load_save_example():
    diff, acc, frozen = _load("tensorflow_load_save").main([])
    assert diff < 1e-4 and acc > 0.8 and frozen


def test_onnx_load_example():
    assert _load("onnx_load").main([]) < 1e-5


def test_compat_example_and_model_modules_import():
    import importlib
    for name in ("bigdl.examples.imageframe.inception_validation", "bigdl.examples.onnx.load_onnx_resnet",
                 "bigdl.examples.keras.keras_utils", "bigdl.examples.lenet.lenet", "bigdl.models.ml_pipeline.dl_classifier",
                 "bigdl.models.utils.model_broadcast", "bigdl.models.textclassifier.textclassifier"):
        importlib.import_module(name)
    from bigdl.models.textclassifier.textclassifier import analyze_texts, build_model, pad
    import torch
    assert analyze_texts([("the cat the dog", 1), ("a cat", 2)])[0] == ("the", (1, 2)) and pad([1], 0, 3) == [1, 0, 0]
    assert tuple(build_model(4, "gru", 16, 12).forward(torch.randn(2, 12, 16)).shape) == (2, 4)
