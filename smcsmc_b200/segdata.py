"""Host-side mirror of the reference's .seg reader (Segment, src/segdata.cpp:55-223,413-451) producing the flat row
arrays the engine uploads.  Same validation rules and error messages; positions are offsets from `data_start`."""
import gzip
from dataclasses import dataclass

import numpy as np


class InvalidSeg(ValueError):
    pass


@dataclass
class SegRows:
    x_start: np.ndarray     # [R] segment_start (offset from data_start)
    x_end: np.ndarray       # [R] min(segment_end, seqlen)
    alleles: np.ndarray     # [R, n] int8: 0, 1, -1 missing, 2 unphased het
    is_site: np.ndarray     # [R] uint8: 1 = SEGMENT_INVARIANT (a typed site ends the row), 0 = INVARIANT_PARTIAL
    dist_to_mutation: np.ndarray   # [R] Segment::distance_to_mutation (segdata.cpp:234-262)
    seqlen: float

    def max_epoch_to_record(self, lags):
        """max_epoch_to_update (smcsmc.cpp:266-275) per row."""
        lags = np.asarray(lags, dtype=np.float64)
        # number of leading epochs e with d < 0.5 * lag[e] (the reference's loop stops at the first that fails), minus one
        ok = np.asarray(self.dist_to_mutation, dtype=np.float64)[:, None] < 0.5 * lags[None, :]
        lead = np.where(ok.all(axis=1), len(lags), np.argmin(ok, axis=1))
        return (lead - 1).astype(np.int32)


_CODE = {".": -1, "/": 2, "0": 0, "1": 1}


def _alleles(field, nsam):
    if nsam > len(field):
        raise InvalidSeg("Number of variant site is wrong: " + field)
    out = []
    for i in range(nsam):
        ch = field[i]
        if ch not in _CODE:
            raise InvalidSeg("Unknown character found in .seg file; expect one of '.', '/', '0' or '1'.")
        v = _CODE[ch]
        out.append(v)
        if v == -1 and i % 2 == 1 and out[i - 1] != -1:
            raise InvalidSeg("Found inconsistent unphased heterozygous marks")
    return out


def parse_seg(path, nsam):
    """Segment::prepare's line parser (segdata.cpp:55-120): every data line of the file, validated, as arrays
    (start[int64], length[int64], alleles[int8, nsam])."""
    opener = gzip.open if str(path).endswith(".gz") else open
    starts, lens, als = [], [], []
    next_start = -1
    with opener(path, "rt") as f:
        for line in f:
            line = line.rstrip("\n")
            if not line:
                break
            if line[0] == "#":
                continue
            cols = line.split("\t")
            if len(cols) < 3:
                raise InvalidSeg("Require 3 or 6 columns")
            try:
                seg_start = int(cols[0])
            except ValueError:
                raise InvalidSeg("Segment start position at:" + line)
            seg_len = int(float(cols[1]))   # the last length may carry a decimal point
            if cols[2] in ("T", "F"):
                if len(cols) != 6:
                    raise InvalidSeg("Require 6 (or 3) columns")
                if cols[3] not in ("T", "F"):
                    raise InvalidSeg("Expected T or F in .seg file column 3 and 4")
                try:
                    int(cols[4])
                except ValueError:
                    raise InvalidSeg("Bad chromosome (not an integer) in column 5")
                al = _alleles(cols[5], nsam)
            else:
                if len(cols) != 3:
                    raise InvalidSeg("Require 3 (or 6) columns")
                al = _alleles(cols[2], nsam)
            if next_start > -1 and next_start != seg_start:
                raise InvalidSeg("Segments are not consecutive")
            next_start = seg_start + seg_len
            starts.append(seg_start); lens.append(seg_len); als.append(al)
    return (np.array(starts, dtype=np.int64), np.array(lens, dtype=np.int64),
            np.array(als, dtype=np.int8).reshape(len(starts), nsam))


def window_rows(raw, seqlen, data_start=1, max_segment_length=1e99, path="<seg>"):
    """The row stream of Segment for the window [data_start, data_start + seqlen): long rows split into
    INVARIANT_PARTIAL pieces (segdata.cpp:125-144), stream cut at the row that reaches seqlen (smcsmc.cpp:353-356)."""
    starts_all, lens_all, al_all = raw
    buf = []   # (start, length, state, allele row index)
    lo = int(np.searchsorted(starts_all + lens_all, data_start, side="right"))
    for idx in range(lo, len(starts_all)):
        seg_start, seg_len = int(starts_all[idx]), int(lens_all[idx])
        next_start = seg_start + seg_len
        if seg_start >= data_start + seqlen:
            break
        if seg_start + seg_len > data_start:
            while True:
                if seg_len > max_segment_length:
                    seg_len = int(max_segment_length)
                    state = 0   # SEGMENT_INVARIANT_PARTIAL
                else:
                    state = 1   # SEGMENT_INVARIANT
                if seg_start + seg_len > data_start:
                    buf.append((seg_start, seg_len, state, idx))
                seg_start += seg_len
                seg_len = next_start - seg_start
                if not seg_start < next_start:
                    break
    if not buf:
        raise InvalidSeg("No data found in file %s between positions %d and %d" % (path, data_start, data_start + seqlen))
    R = len(buf)
    starts = np.array([b[0] for b in buf], dtype=np.int64)
    lens = np.array([b[1] for b in buf], dtype=np.int64)
    al = al_all[np.array([b[3] for b in buf])]
    state = np.array([b[2] for b in buf], dtype=np.uint8)
    s = np.maximum(starts - data_start, 0).astype(np.float64)
    e = (starts - data_start + lens).astype(np.float64)
    stop = R
    hit = np.nonzero(e >= seqlen)[0]
    if len(hit):
        stop = int(hit[0]) + 1
    allmiss = (al == -1).all(axis=1)
    dist = np.zeros(R)
    for r in np.nonzero(allmiss)[0]:
        i = r
        d = 0.0
        while i >= 0 and allmiss[i]:
            d = float(starts[r] - starts[i])
            i -= 1
        j = r
        while j < R and allmiss[j]:
            j += 1
        if j < R:
            d = min(d, float(starts[j] + lens[j] - starts[r]))
        dist[r] = d
    return SegRows(s[:stop], np.minimum(e[:stop], seqlen), al[:stop], state[:stop], dist[:stop], float(seqlen))


def read_seg(path, nsam, seqlen, data_start=1, max_segment_length=1e99):
    """Segment::prepare + the row stream of read_new_line, up to the row that reaches seqlen."""
    return window_rows(parse_seg(path, nsam), seqlen, data_start, max_segment_length, path)
