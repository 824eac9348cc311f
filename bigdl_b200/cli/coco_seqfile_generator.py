"""COCO images + annotation json → SequenceFiles (S/models/utils/COCOSeqFileGenerator.scala): key = image file name,
value = encoded image bytes followed by a compact annotation block (boxes, category indices, iscrowd flags, RLE masks)."""
from __future__ import annotations

import argparse
import json
import os
import struct


def main(argv=None):
    ap = argparse.ArgumentParser(description="COCO Sequence File Generator")
    ap.add_argument("-f", "--folder", required=True, help="image folder")
    ap.add_argument("-m", "--metaPath", required=True, help="instances_*.json")
    ap.add_argument("-o", "--output", required=True)
    ap.add_argument("-b", "--blockSize", type=int, default=12800)
    a = ap.parse_args(argv)
    from ..dataset.segmentation import COCODataset, MaskUtils
    from ..dataset.seqfile import SequenceFileWriter

    ds = COCODataset.load(a.metaPath, a.folder)
    os.makedirs(a.output, exist_ok=True)
    writer, n, idx = None, 0, 0
    for im in ds.images:
        if writer is None:
            writer = SequenceFileWriter(os.path.join(a.output, f"coco-seq-{idx}.seq"))
        with open(im.file_name, "rb") as f:
            img = f.read()
        anns = []
        for an in im.annotations:
            rle = an.segmentation.toRLE() if an.segmentation is not None else None
            anns.append({"bbox": list(an.bbox), "category": ds.categoryId2Idx[an.category_id], "iscrowd": int(an.iscrowd),
                         "rle": MaskUtils.RLE2String(rle) if rle is not None else None})
        meta = json.dumps({"height": im.height, "width": im.width, "annotations": anns}).encode()
        writer.append(os.path.basename(im.file_name).encode(), struct.pack(">i", len(img)) + img + meta)
        n += 1
        if n == a.blockSize:
            writer.close()
            writer, n, idx = None, 0, idx + 1
    if writer is not None:
        writer.close()
    print(f"{len(ds.images)} images → {idx + (1 if n else 0)} sequence files under {a.output}")


if __name__ == "__main__":
    main()
