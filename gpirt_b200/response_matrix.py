"""Data front-end of the sampler path: the reference's datasets and its response coding, on the host.

* read_rda()            - reader for the `.rda` files the reference ships (data/senate116.rda: RDX2 + bzip2, data/SDO.rda: RDX3 + gzip;
                          R's XDR serialisation), enough of the format for data frames of integer / double / string / factor columns.
* spread()              - the long -> wide reshape of the vignette (vignettes/gpirt-vignette.Rmd:135-144: icpsr x rollnumber -> cast_code).
* response_matrix()     - R/response_matrix.R:51-99: yea -> 1, nay -> -1, anything else -> NA (with the message), unanimous items dropped.
* is_response_matrix(), as_response_matrix() - R/response_matrix.R:109-127.
* ordinal_codes()       - what the C++ sampler actually consumes (src/log-likelihood.cpp:57: c = int(y) indexes thresholds[c-1], thresholds[c]):
                          an n x m x horizon cube of codes 1..C with NaN for missing. The reference's own recoding path hands it {-1, 1, NA}
                          (that path is stale and indexes out of range); binary data are C = 2 with nay -> 1, yea -> 2.
* posterior summaries   - theta_mean() (the vignette's "mean of the columns of samples[['theta']]") and irf_mean() (expected response curves
                          over the theta grid from stored f* draws, the quantity the reference's plots are made of).

Nothing here touches the GPU; it feeds gpirt_b200.gpirtMCMC().
"""
import bz2
import gzip
import lzma
import struct

import numpy as np

NA_INTEGER = -2147483648


class RdaError(ValueError):
    pass


class _Reader:
    def __init__(self, buf):
        self.b, self.p, self.refs = buf, 0, []

    def int(self):
        v = struct.unpack_from(">i", self.b, self.p)[0]
        self.p += 4
        return v

    def length(self):
        n = self.int()
        if n == -1:  # long vector: two 32-bit halves
            hi, lo = self.int(), self.int()
            n = (hi << 32) + (lo & 0xFFFFFFFF)
        return n

    def bytes(self, n):
        v = self.b[self.p:self.p + n]
        self.p += n
        return v

    def item(self):
        flags = self.int()
        t, has_attr, has_tag = flags & 0xFF, bool(flags & 0x200), bool(flags & 0x400)
        if t == 254:  # NILVALUE
            return None
        if t in (253, 252, 251, 250, 242, 241):  # global / missing-arg / unbound / empty / base environments
            return ("env", t)
        if t == 255:  # REFSXP
            idx = flags >> 8
            if idx == 0:
                idx = self.int()
            return self.refs[idx - 1]
        if t == 1:  # SYMSXP
            name = self.item()
            self.refs.append(name)
            return name
        if t in (2, 6):  # LISTSXP / LANGSXP -> python list of (tag, value); attributes dropped
            out = []
            while True:
                if has_attr:
                    self.item()
                tag = self.item() if has_tag else None
                out.append((tag, self.item()))
                flags = self.int()
                t2 = flags & 0xFF
                if t2 == 254:
                    return out
                if t2 not in (2, 6):
                    raise RdaError(f"pairlist tail of type {t2}")
                has_attr, has_tag = bool(flags & 0x200), bool(flags & 0x400)
        if t == 9:  # CHARSXP
            n = self.int()
            if n == -1:
                return None
            return self.bytes(n).decode("utf-8", "replace")
        if t == 238:  # ALTREP: (info pairlist, state, attributes)
            info, state, attr = self.item(), self.item(), self.item()
            cls = info[0][1]
            if cls == "compact_intseq":
                n, start, step = (float(v) for v in state["value"])
                value = dict(value=(start + step * np.arange(int(n))).astype(np.int32), attr={})
            elif cls == "compact_realseq":
                n, start, step = (float(v) for v in state["value"])
                value = dict(value=start + step * np.arange(int(n)), attr={})
            elif cls.startswith("wrap_"):
                value = state["value"][0]
            elif cls == "deferred_string":
                src = state[0][1]
                value = dict(value=[None if _is_na(v) else _fmt(v) for v in src["value"]], attr={})
            else:
                raise RdaError(f"ALTREP class {cls} not handled")
            if attr:
                value = dict(value=value["value"], attr={**value["attr"], **dict(attr)})
            return value
        if t in (10, 13):  # LGLSXP, INTSXP
            n = self.length()
            v = np.frombuffer(self.bytes(4 * n), dtype=">i4").astype(np.int32)
        elif t == 14:  # REALSXP
            n = self.length()
            v = np.frombuffer(self.bytes(8 * n), dtype=">f8").astype(np.float64)
        elif t == 16:  # STRSXP
            n = self.length()
            v = [self.item() for _ in range(n)]
        elif t == 19:  # VECSXP
            n = self.length()
            v = [self.item() for _ in range(n)]
        else:
            raise RdaError(f"SEXP type {t} not handled at byte {self.p}")
        attr = dict(self.item() or []) if has_attr else {}
        return dict(value=v, attr=attr)


def _is_na(v):
    return v is None or (isinstance(v, (int, np.integer)) and v == NA_INTEGER) or (isinstance(v, float) and np.isnan(v))


def _fmt(v):
    return str(int(v)) if float(v) == int(v) else repr(float(v))


def _column(obj):
    """One data-frame column as float64 (NA -> NaN), or a list of strings (factor levels applied)."""
    v, attr = obj["value"], obj["attr"]
    if "levels" in attr:  # factor: integer codes into the level strings
        lev = attr["levels"]["value"]
        return [None if c == NA_INTEGER else lev[c - 1] for c in v]
    if isinstance(v, list):
        return v
    if v.dtype == np.int32:
        out = v.astype(np.float64)
        out[v == NA_INTEGER] = np.nan
        return out
    return v


def read_rda(path):
    """{object name: value}; a data frame becomes {column name: float64 array (NA = NaN) or list of str}."""
    raw = open(path, "rb").read()
    for opener in (bz2.decompress, gzip.decompress, lzma.decompress, lambda b: b):
        try:
            buf = opener(raw)
            break
        except Exception:
            continue
    if buf[:5] not in (b"RDX2\n", b"RDX3\n") or buf[5:7] != b"X\n":
        raise RdaError("not an XDR .rda file")
    r = _Reader(buf)
    r.p = 7
    version = r.int()
    r.int()
    r.int()
    if version == 3:
        r.bytes(r.int())  # native encoding name
    out = {}
    for name, obj in r.item():
        if isinstance(obj, dict) and isinstance(obj["value"], list) and "names" in obj["attr"]:
            out[name] = {cn: _column(col) for cn, col in zip(obj["attr"]["names"]["value"], obj["value"])}
        else:
            out[name] = obj["value"] if isinstance(obj, dict) else obj
    return out


def spread(frame, row_key, col_key, value):
    """tidyr::spread(col_key, value) after select(row_key, col_key, value): rows and columns in ascending key order, absent cells NaN.
    Returns (matrix, row keys, column keys)."""
    rk, ck, v = (np.asarray(frame[k], dtype=np.float64) for k in (row_key, col_key, value))
    rows, ri = np.unique(rk, return_inverse=True)
    cols, ci = np.unique(ck, return_inverse=True)
    out = np.full((len(rows), len(cols)), np.nan)
    out[ri, ci] = v
    return out, rows, cols


DEFAULT_CODES = dict(yea=(1, 2, 3), nay=(4, 5, 6), missing=(0, 7, 8, 9, None))


def _is_na_code(c):
    return c is None or (isinstance(c, float) and np.isnan(c))


def _code_list(codes, key):
    c = codes[key]
    return list(c) if isinstance(c, (list, tuple, range, np.ndarray)) else [c]


def response_matrix(data, response_codes=None, messages=None):
    """R/response_matrix.R:51-99. `data`: an n x m array of numbers (NaN = NA) or of strings / None, or a data frame given as
    {column name: sequence} (factor columns as their level strings). A bare Python list is refused like an R list that is not a data
    frame (:55-58). Returns (matrix in {1, -1, NaN} with the unanimous items removed, kept column indices). `messages`, if a list,
    receives the messages the reference prints (unlisted response values, discarded unanimous items)."""
    codes = DEFAULT_CODES if response_codes is None else response_codes
    if isinstance(data, list):
        raise ValueError("Conversion from lists to response_matrix objects is currently unsupported.")
    if isinstance(data, dict):  # data frame -> as.matrix (:65)
        cols = [np.asarray(c, dtype=object) if isinstance(c, list) else np.asarray(c) for c in data.values()]
        numeric = all(c.dtype.kind in "fiub" for c in cols)
        tmp = np.stack([c.astype(np.float64) if numeric else c.astype(object) for c in cols], axis=1)
    else:
        tmp = np.asarray(data)
    if tmp.ndim != 2:
        raise ValueError("response_matrix: data must be a matrix or a data frame")
    if tmp.dtype.kind in "fiub":
        tmp = tmp.astype(np.float64)
        na = np.isnan(tmp)

        def members(key):
            vals = [float(c) for c in _code_list(codes, key) if not _is_na_code(c) and not isinstance(c, str)]
            return np.isin(tmp, np.asarray(vals, dtype=np.float64)) if vals else np.zeros(tmp.shape, bool)
    else:
        tmp = tmp.astype(object)
        na = np.frompyfunc(_is_na_code, 1, 1)(tmp).astype(bool)

        def members(key):
            vals = {c for c in _code_list(codes, key) if not _is_na_code(c)}
            return np.frompyfunc(lambda v: (not _is_na_code(v)) and v in vals, 1, 1)(tmp).astype(bool)

    yea, nay = members("yea"), members("nay")
    miss = members("missing") | (na if any(_is_na_code(c) for c in _code_list(codes, "missing")) else np.zeros(tmp.shape, bool))
    surprise = ~(yea | nay | miss)
    if surprise.any() and messages is not None:  # :69-75 - unlisted values are treated as missing, with a message
        vals = sorted({("NA" if _is_na_code(v) else (v if isinstance(v, str) else _fmt(v))) for v in tmp[surprise]})
        messages.append("Responses with value " + ", ".join(vals) + " were not given a response code and will be treated as missing.")
    result = np.full(tmp.shape, np.nan)
    result[yea] = 1.0  # :77-79, in this order (a code listed twice ends up as the later assignment)
    result[nay] = -1.0
    result[miss | surprise] = np.nan
    # :85-88 - an item is unanimous when its non-missing responses take exactly one value (an all-NA item is NOT dropped: length 0)
    unanimous = np.array([len(np.unique(col[~np.isnan(col)])) == 1 for col in result.T], dtype=bool)
    if unanimous.any() and messages is not None:
        idx = np.flatnonzero(unanimous) + 1
        messages.append(("Items " if len(idx) > 1 else "Item ") + ", ".join(str(i) for i in idx) + (" were" if len(idx) > 1 else " was") +
                        " discarded as unanimous.")
    return np.asfortranarray(result[:, ~unanimous]), np.flatnonzero(~unanimous)


def is_response_matrix(x):
    """R/response_matrix.R:109-115 (class tag aside): a matrix with values in {NA, -1, 1} only."""
    x = np.asarray(x)
    return x.ndim == 2 and x.dtype.kind == "f" and bool(np.all(np.isnan(x) | (x == 1.0) | (x == -1.0)))


def as_response_matrix(x, response_codes=None, messages=None):
    """R/response_matrix.R:119-127."""
    if is_response_matrix(x):
        return np.asfortranarray(x)
    return response_matrix(x, response_codes, messages)[0]


def ordinal_codes(y):
    """Codes 1..C (NaN missing) as an n x m x horizon cube for the sampler: the sorted distinct observed values are ranked 1..C
    (R/gpirtMCMC.R:145 takes C = number of distinct observed values). A {-1, 1, NA} response_matrix becomes nay -> 1, yea -> 2."""
    y = np.asarray(y, dtype=np.float64)
    if y.ndim == 2:
        y = y[:, :, None]
    vals = np.unique(y[np.isfinite(y)])
    out = np.full(y.shape, np.nan)
    for rank, v in enumerate(vals):
        out[y == v] = rank + 1.0
    return np.asfortranarray(out), len(vals)


def senate116_responses(path):
    """BASELINE config 1: data/senate116.rda -> spread -> response_matrix -> codes {1 (nay), 2 (yea)}: (n x m x 1 cube, C, senator ids, rollcalls)."""
    df = read_rda(path)["senate116"]
    wide, icpsr, rollnumber = spread(df, "icpsr", "rollnumber", "cast_code")
    rm, kept = response_matrix(wide)
    cube, C = ordinal_codes(rm)
    return cube, C, icpsr, rollnumber[kept]


def sdo_responses(path):
    """BASELINE config 2: data/SDO.rda, the Likert items as an n x m x 1 cube of codes 1..C."""
    objs = read_rda(path)
    name = "SDO" if "SDO" in objs else next(iter(objs))
    df = objs[name]
    if isinstance(df, dict):
        cols = [np.asarray(c, dtype=np.float64) for c in df.values() if not isinstance(c, list)]
        mat = np.stack(cols, axis=1)
    else:
        mat = np.asarray(df, dtype=np.float64)
    cube, C = ordinal_codes(mat)
    return cube, C


# ---- posterior summaries ----

def theta_mean(samples):
    """Posterior mean of theta per respondent and session: colMeans of samples[['theta']] (vignette, last paragraph)."""
    return np.asarray(samples["theta"]).mean(axis=0)


def irf_mean(fstar_draws, beta_draws, threshold_draws, grid=None):
    """Posterior mean expected response E[y | theta*] on the grid for every item (slice 0): for each stored draw,
    g = f* + mu* with mu* = (1, theta*, theta*^2) beta (src/gpirtMCMC.cpp:308-311) and P(y = c) = Phi(tau_c - g) - Phi(tau_{c-1} - g)
    (src/log-likelihood.cpp:57-62, without the 1e-6 floor, which only guards the log)."""
    from scipy.special import ndtr

    grid = np.arange(-5.0, 5.0 + 1e-9, 0.01)[:1001] if grid is None else np.asarray(grid)
    X = np.stack([np.ones_like(grid), grid, grid * grid], axis=1)
    acc = 0.0
    for fs, be, th in zip(fstar_draws, beta_draws, threshold_draws):
        fs, be, th = (np.asarray(a)[..., 0] if np.asarray(a).ndim == 3 else np.asarray(a) for a in (fs, be, th))
        g = fs + X @ be  # N x m
        Ccat = th.shape[1] - 1
        cdf = ndtr(th.T[:, None, :] - g[None, :, :])  # (C+1) x N x m
        p = cdf[1:] - cdf[:-1]
        acc = acc + np.tensordot(np.arange(1, Ccat + 1, dtype=np.float64), p, axes=(0, 0))
    return acc / len(fstar_draws)
