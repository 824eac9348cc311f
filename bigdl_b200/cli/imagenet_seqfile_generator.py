"""ImageNet folder → Hadoop SequenceFiles (S/models/utils/ImageNetSeqFileGenerator.scala): ``-f <imagenet root with
train/ and val/ class folders> -o <output> [-p parallel] [-b blockSize] [-r] [--scaleSize 256]``. Each record is
key = "name\\nlabel" (or "label"), value = 4-byte BE width, height + BGR bytes."""
from __future__ import annotations

import argparse
import os
from concurrent.futures import ThreadPoolExecutor


def _convert(paths, base, block_size, has_name, scale_to):
    from ..dataset.image import read_image_bgr
    from ..dataset.seqfile import write_image_seq_files

    def items():
        for p in paths:
            yield read_image_bgr(p.path, scale_to), int(p.label), os.path.basename(p.path)
    return write_image_seq_files(items(), base, block_size, has_name)


def main(argv=None):
    ap = argparse.ArgumentParser(description="Imagenet Sequence File Generator")
    ap.add_argument("-f", "--folder", required=True)
    ap.add_argument("-o", "--output", required=True)
    ap.add_argument("-p", "--parallel", type=int, default=1)
    ap.add_argument("-b", "--blockSize", type=int, default=12800)
    ap.add_argument("-t", "--trainOnly", action="store_true")
    ap.add_argument("-v", "--validationOnly", action="store_true")
    ap.add_argument("-r", "--resize", action="store_true", help="kept for CLI parity: images are always scaled")
    ap.add_argument("-n", "--hasName", action="store_true")
    ap.add_argument("--scaleSize", type=int, default=256)
    a = ap.parse_args(argv)
    from ..dataset.image import LocalImageFiles

    for split, skip in (("train", a.validationOnly), ("val", a.trainOnly)):
        src = os.path.join(a.folder, split)
        if skip or not os.path.isdir(src):
            continue
        out_dir = os.path.join(a.output, split)
        os.makedirs(out_dir, exist_ok=True)
        paths = LocalImageFiles.readPaths(src)
        shards = [paths[i::a.parallel] for i in range(a.parallel)]
        with ThreadPoolExecutor(max_workers=a.parallel) as ex:
            futs = [ex.submit(_convert, sh, os.path.join(out_dir, f"imagenet-seq-{i}"), a.blockSize, a.hasName, a.scaleSize)
                    for i, sh in enumerate(shards) if sh]
            files = [f for fu in futs for f in fu.result()]
        print(f"{split}: {len(paths)} images → {len(files)} sequence files under {out_dir}")


if __name__ == "__main__":
    main()
