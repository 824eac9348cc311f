from __future__ import annotations

import os

import numpy as np

from ..transform.vision.image import BytesToMat, FeatureTransformer, ImageFeature


class DLImageReader:
    """``readImages(path)`` → DataFrame with one ``image`` column of records (origin, height, width, nChannels, mode,
    data) — the schema of DLImageReader.scala:40-110 (``mode`` is the OpenCV type code, data = raw BGR bytes)."""

    @staticmethod
    def readImages(path, sc=None, minPartitions=1):
        import pandas as pd

        paths = [path] if os.path.isfile(path) else sorted(os.path.join(path, f) for f in os.listdir(path)
                                                           if os.path.isfile(os.path.join(path, f)))
        rows = []
        for p in paths:
            with open(p, "rb") as f:
                feat = BytesToMat().transform(ImageFeature(f.read(), path=p))
            if not feat.isValid:
                continue
            m = feat["mat"]
            rows.append({"origin": p, "height": m.shape[0], "width": m.shape[1], "nChannels": m.shape[2],
                         "mode": 16 if m.shape[2] == 3 else 0, "data": m.tobytes()})
        return pd.DataFrame({"image": rows})


class DLImageTransformer:
    """Apply a vision ``FeatureTransformer`` to an image column (DLImageTransformer.scala:40-120); the output column
    holds float records (mode 21 = CV_32FC3) like the reference."""

    def __init__(self, transformer: FeatureTransformer):
        self.transformer = transformer
        self.inputCol, self.outputCol = "image", "output"

    def setInputCol(self, v): self.inputCol = v; return self
    def setOutputCol(self, v): self.outputCol = v; return self

    def transform(self, df):
        out_rows = []
        for rec in df[self.inputCol]:
            mat = np.frombuffer(rec["data"], dtype=np.uint8 if rec["mode"] in (0, 16) else np.float32)
            mat = mat.reshape(rec["height"], rec["width"], rec["nChannels"]).copy()
            f = self.transformer.transform(ImageFeature(mat, path=rec["origin"]))
            m = np.asarray(f["mat"], dtype=np.float32)
            out_rows.append({"origin": rec["origin"], "height": m.shape[0], "width": m.shape[1], "nChannels": m.shape[2],
                             "mode": 21 if m.shape[2] == 3 else 5, "data": m.tobytes()})
        out = df.copy()
        out[self.outputCol] = out_rows
        return out
