"""Stand-in for the ``pyBigWig`` module -- TEST INFRASTRUCTURE ONLY (oracle side).

The compiled reference's ``PyMaSC.reader.bigwig`` does ``import pyBigWig`` at import
time (reader/bigwig.pyx:19) and uses only ``open()``, ``.chroms()``, ``.intervals()``
and ``.close()`` (:140-171, :188-198).  pyBigWig is not installable here, so this
module answers those four calls from the text twin of a track: ``X.bigwig`` is served
from ``X.bedGraph`` beside it, with chromosome sizes from any ``*.chrom.sizes`` in the
same directory.  It never decodes the binary BigWig format.
"""
import glob
import io
import os


class _Track(object):
    def __init__(self, path):
        stem, _ = os.path.splitext(path)
        text = stem + ".bedGraph"
        if not os.path.exists(text):
            raise RuntimeError("pyBigWig stub: no bedGraph twin for %s" % path)
        self._by_chrom = {}
        with io.open(text) as fh:      # (this module's own open() below is pyBigWig's)
            for line in fh:
                if not line.strip() or line.startswith(("#", "track", "browser")):
                    continue
                chrom, start, end, value = line.split()[:4]
                self._by_chrom.setdefault(chrom, []).append((int(start), int(end), float(value)))
        self._sizes = {}
        for sizes in glob.glob(os.path.join(os.path.dirname(path) or ".", "*.chrom.sizes")):
            with io.open(sizes) as fh:
                for line in fh:
                    parts = line.split()
                    if len(parts) >= 2:
                        self._sizes[parts[0]] = int(parts[1])
        for chrom, ivs in self._by_chrom.items():
            self._sizes.setdefault(chrom, max(e for _, e, _ in ivs))

    def chroms(self):
        return dict(self._sizes)

    def intervals(self, chrom, start=0, end=None):
        ivs = self._by_chrom.get(chrom)
        if ivs is None:
            return None
        return [iv for iv in ivs if iv[1] > start and (end is None or iv[0] < end)]

    def close(self):
        self._by_chrom = {}


def open(path, mode="r"):  # noqa: A001 - pyBigWig's own name
    return _Track(path)
