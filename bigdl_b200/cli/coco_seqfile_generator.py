"""COCO images + annotation json -> SequenceFiles (S/models/utils/COCOSeqFileGenerator.scala:71-116).

Record layout of the reference: key = ``COCOSerializeContext`` bytes (file name, height, width, annotations with
category index / area / bbox / iscrowd / polygons or RLE counts, closed by ``COCODataset.MAGIC_NUM``), value = the raw
image file bytes. ``read_coco_seqfile`` is the matching reader (DataSet.SeqFileFolder.filesToRoiImageFeatures)."""
from __future__ import annotations

import argparse
import os


def main(argv=None):
    ap = argparse.ArgumentParser(description="COCO Sequence File Generator")
    ap.add_argument("-f", "--folder", required=True, help="image folder")
    ap.add_argument("-m", "--metaPath", required=True, help="instances_*.json")
    ap.add_argument("-o", "--output", required=True)
    ap.add_argument("-p", "--parallel", type=int, default=1, help="accepted for compatibility; blocks are written in order")
    ap.add_argument("-b", "--blockSize", type=int, default=12800)
    a = ap.parse_args(argv)
    from ..dataset.segmentation import COCODataset
    from ..dataset.seqfile import SequenceFileWriter

    ds = COCODataset.load(a.metaPath, a.folder)
    os.makedirs(a.output, exist_ok=True)
    images = []
    for im in ds.images:
        if os.path.isfile(im.path):
            images.append(im)
        else:
            print(f"[Warning] The image file {os.path.basename(im.path)} does not exist.")
    blocks = 0
    for blk, start in enumerate(range(0, len(images), a.blockSize)):
        writer = SequenceFileWriter(os.path.join(a.output, f"coco-seq-{blk}.seq"))
        for im in images[start:start + a.blockSize]:
            writer.append(ds.imageKey(im), im.data)
        writer.close()
        blocks += 1
    print(f"{len(images)} images -> {blocks} sequence files under {a.output}")


def read_coco_seqfile(path: str):
    """Yield ``(file_name, height, width, [SimpleAnnotation], image_bytes)`` from one generated file."""
    from ..dataset.segmentation import COCODataset
    from ..dataset.seqfile import read_sequence_file

    for key, value in read_sequence_file(path):
        name, h, w, anns = COCODataset.parseKey(key)
        yield name, h, w, anns, value


if __name__ == "__main__":
    main()
