from bigdl_b200.dataset.base import *  # noqa: F401,F403

import os as _os
import sys as _sys
import time as _time


class Progbar:
    """Console progress bar of the reference's download helper (P/dataset/base.py:28-133)."""

    def __init__(self, target, width=30, verbose=1, interval=0.01):
        self.target, self.width, self.verbose, self.interval = target, width, verbose, interval
        self.sum_values, self.unique_values = {}, []
        self.start, self.last_update, self.total_width, self.seen_so_far = _time.time(), 0, 0, 0

    def update(self, current, values=None, force=False):
        for k, v in (values or []):
            if k not in self.sum_values:
                self.sum_values[k] = [v * (current - self.seen_so_far), current - self.seen_so_far]
                self.unique_values.append(k)
            else:
                self.sum_values[k][0] += v * (current - self.seen_so_far)
                self.sum_values[k][1] += current - self.seen_so_far
        self.seen_so_far = current
        now = _time.time()
        if not self.verbose or (not force and now - self.last_update < self.interval):
            return
        self.last_update = now
        frac = min(float(current) / self.target, 1.0) if self.target else 1.0
        done = int(self.width * frac)
        bar = "[" + "=" * max(done - 1, 0) + (">" if 0 < done < self.width else "=" * min(done, 1)) + "." * (self.width - done) + "]"
        info = "".join(" - %s: %.4f" % (k, self.sum_values[k][0] / max(1, self.sum_values[k][1])) for k in self.unique_values)
        _sys.stdout.write("\r%d/%d %s%s" % (current, self.target, bar, info))
        if current >= self.target:
            _sys.stdout.write("\n")
        _sys.stdout.flush()

    def add(self, n, values=None):
        self.update(self.seen_so_far + n, values)


def display_table(rows, positions):
    for objects in rows:
        line = ""
        for obj, pos in zip(objects, positions):
            line = (line + str(obj))[:pos]
            line += " " * (pos - len(line))
        print(line)


def maybe_download(filename, work_directory, source_url):
    """Return the local path of ``filename`` under ``work_directory``, fetching it from ``source_url`` when absent.
    Machines without network access get a clear error naming the file to place there."""
    _os.makedirs(work_directory, exist_ok=True)
    filepath = _os.path.join(work_directory, filename)
    if _os.path.exists(filepath):
        return filepath
    try:
        from urllib.request import urlretrieve
        bar = Progbar(1)

        def hook(count, block, total):
            bar.target = max(total, 1)
            bar.update(min(count * block, bar.target))
        urlretrieve(source_url, filepath, hook)
    except Exception as e:
        if _os.path.exists(filepath):
            _os.remove(filepath)
        raise IOError(f"cannot download {source_url} ({e}); place the file at {filepath}") from e
    return filepath
