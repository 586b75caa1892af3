"""TEST DOUBLE for libxchg_b200.so -- host-memory numpy stand-in for the C-ABI.

Lets the ``-m "not gpu"`` suite exercise the Python host layer (Arrow staging,
boundary normalisation, partition slicing, export, the exchange schedule) in a
container without a GPU: pointers are host pointers (CPU torch tensors) and each
entry point is an independent numpy restatement of the kernel's contract in
include/xchg_b200.h.  It is installed only by tests (``_cabi._install_test_double``);
the product path never sees it.
"""
from __future__ import annotations

import ctypes

import numpy as np


def _addr(p):
    if p is None:
        return 0
    if isinstance(p, ctypes.c_void_p):
        return p.value or 0
    if isinstance(p, int):
        return p
    return ctypes.addressof(p)


def _arr(p, dtype, count):
    a = _addr(p)
    dt = np.dtype(dtype)
    if count == 0:
        return np.zeros(0, dtype=dt)
    assert a != 0, "NULL pointer"
    buf = (ctypes.c_uint8 * (count * dt.itemsize)).from_address(a)
    return np.frombuffer(buf, dtype=dt, count=count)


_KT = {0: np.int8, 1: np.int16, 2: np.int32, 3: np.int64, 4: np.uint8, 5: np.uint16,
       6: np.uint32, 7: np.uint64, 8: np.float32, 9: np.float64}


def _bits(p, bit_off, idx):
    a = _addr(p)
    pos = idx.astype(np.int64) + bit_off
    nbytes = int(pos.max()) // 8 + 1 if len(pos) else 0
    bm = _arr(p, np.uint8, nbytes)
    return ((bm[pos >> 3] >> (pos & 7).astype(np.uint8)) & 1).astype(bool)


def _normalize(vals, kt, desc):
    dt = np.dtype(_KT[kt])
    bits = dt.itemsize * 8
    mask = np.uint64((1 << bits) - 1) if bits < 64 else np.uint64(0xFFFFFFFFFFFFFFFF)
    if dt.kind == "f":
        u = np.uint32 if bits == 32 else np.uint64
        v = vals.copy()
        v[v == 0] = 0.0
        b = v.view(u).astype(np.uint64)
        sign = np.uint64(1 << (bits - 1))
        t = np.where(b & sign != 0, ~b & mask, b | sign)
        if desc:
            t = ~t & mask
        t = np.where(np.isnan(vals), np.uint64(int(mask) - 1), t)
        return t.astype(np.uint64)
    if dt.kind == "i":
        u = (vals.astype(np.int64).view(np.uint64) + np.uint64(1 << (bits - 1))) & mask
    else:
        u = vals.astype(np.uint64)
    if desc:
        u = ~u & mask
    return u.astype(np.uint64)


def _null_key(kt):
    bits = np.dtype(_KT[kt]).itemsize * 8
    return np.uint64(0xFFFFFFFFFFFFFFFF) if bits == 64 else np.uint64(1 << bits)


class FakeLib:
    def __init__(self):
        self.err = b""
        self.calls = []

    def xchg_abi_version(self):
        return 1

    def xchg_last_error(self):
        return self.err

    def xchg_device_count(self):
        return 0

    def xchg_sort_workspace_bytes(self, n, out):
        out._obj.value = 64
        return 0

    def _norm_col(self, keys, kt, desc, validity, vbit, perm, n):
        dt = _KT[kt]
        rows = _arr(perm, np.uint32, n).astype(np.int64) if _addr(perm) else np.arange(n)
        nrows_src = int(rows.max()) + 1 if n else 0
        src = _arr(keys, dt, nrows_src)
        nk = _normalize(src[rows], kt, bool(desc))
        valid = np.ones(n, dtype=bool)
        if _addr(validity):
            valid = _bits(validity, vbit, rows)
            nk = np.where(valid, nk, _null_key(kt))
        return rows, nk, valid

    def xchg_sort_column(self, keys, kt, desc, validity, vbit, perm_in, n, perm_out, norm_out, ws, wsb, stream):
        self.calls.append("sort_column")
        if n == 0:
            return 0
        rows, nk, valid = self._norm_col(keys, kt, desc, validity, vbit, perm_in, n)
        order = np.lexsort((nk, ~valid))  # stable: valid first, then by key
        order = np.argsort(nk, kind="stable")
        order = order[np.argsort(~valid[order], kind="stable")]
        _arr(perm_out, np.uint32, n)[:] = rows[order].astype(np.uint32)
        if _addr(norm_out):
            _arr(norm_out, np.uint64, n)[:] = nk[order]
        return 0

    def xchg_normalize_keys(self, keys, kt, desc, validity, vbit, perm, n, out, stream):
        if n == 0:
            return 0
        _, nk, _ = self._norm_col(keys, kt, desc, validity, vbit, perm, n)
        _arr(out, np.uint64, n)[:] = nk
        return 0

    def xchg_set_l2_fetch_granularity(self, b):
        return 0

    def xchg_denormalize_keys(self, norm, kt, desc, n, out, stream):
        dt = np.dtype(_KT[kt])
        assert dt.kind in "iu"
        bits = dt.itemsize * 8
        mask = (1 << bits) - 1
        u = _arr(norm, np.uint64, n).astype(object)
        res = []
        for x in u:
            x = int(x) & mask
            if desc:
                x = ~x & mask
            if dt.kind == "i":
                x ^= 1 << (bits - 1)
                if x >= 1 << (bits - 1):
                    x -= 1 << bits
            res.append(x)
        _arr(out, dt, n)[:] = np.array(res, dtype=dt)
        return 0

    def xchg_partition_bounds(self, cols, col_desc, last_thr, has_last, ncols, n, lower, upper, flags,
                              nb, first_desc, cuts, stream):
        lo_a = _arr(lower, np.uint64, nb * ncols)
        up_a = _arr(upper, np.uint64, nb * ncols)
        fl_a = _arr(flags, np.uint8, nb * ncols) if _addr(flags) else np.zeros(nb * ncols, np.uint8)
        out = _arr(cuts, np.uint64, nb)
        arrays = [_arr(cols[c], np.uint64, n) for c in range(ncols)]
        cdesc = [col_desc[c] for c in range(ncols)]
        thr = [last_thr[c] for c in range(ncols)]
        hl = [has_last[c] for c in range(ncols)]
        for b in range(nb):
            lo, hi = 0, n
            early = False
            for c in range(ncols):
                if lo == hi:
                    early = True
                    break
                f = int(fl_a[b * ncols + c])
                if f & 2:
                    break
                if f & 4:
                    if cdesc[c]:
                        hi = lo
                    else:
                        lo = hi
                    continue
                L, U = int(lo_a[b * ncols + c]), int(up_a[b * ncols + c])
                if not cdesc[c]:
                    seg = arrays[c][lo:hi] if hi > lo else arrays[c][0:0]
                    l2 = lo + int(np.searchsorted(seg, np.uint64(L), side="left"))
                    h2 = l2 if (f & 1) else lo + int(np.searchsorted(seg, np.uint64(U), side="right"))
                    lo, hi = l2, h2
                    continue
                # descending: numpy.searchsorted over the reversed view, probe by probe
                a = arrays[c]
                m = hi - lo if hi > lo else 0
                key_last = bool(f & 8)

                def is_last(k):
                    return bool(hl[c]) and k >= thr[c]

                mn, mx = 0, m
                while mn < mx:
                    mid = mn + ((mx - mn) >> 1)
                    k = int(a[hi - 1 - mid])
                    el = is_last(k)
                    before_eq = (k < L) if (f & 1) else (k <= U)
                    lt = (not el) if key_last else ((not el) and (not before_eq))
                    if lt:
                        mn = mid + 1
                    else:
                        mx = mid
                left_idx = mn
                mn, mx = 0, m
                while mn < mx:
                    mid = mn + ((mx - mn) >> 1)
                    k = int(a[hi - 1 - mid])
                    el = is_last(k)
                    key_lt_arr = (not key_last) and (el or k < L)
                    if not key_lt_arr:
                        mn = mid + 1
                    else:
                        mx = mid
                right_idx = mn
                base = lo
                lo, hi = base + (m - right_idx), base + (m - left_idx)
            out[b] = hi if (early or first_desc) else lo
        return 0

    def xchg_gather_fixed(self, inp, width, idx, n, out, stream):
        if n == 0:
            return 0
        ix = _arr(idx, np.uint32, n).astype(np.int64)
        src = _arr(inp, np.uint8, (int(ix.max()) + 1) * width).reshape(-1, width)
        _arr(out, np.uint8, n * width).reshape(-1, width)[:] = src[ix]
        return 0

    def xchg_gather_rows(self, inp, width, idx, n, out, stream):
        return self.xchg_gather_fixed(inp, width, idx, n, out, stream)

    def xchg_gather_rows_multi(self, inp, width, idx, nseg, row0, rows, outs, stream):
        for d in range(nseg):
            if rows[d]:
                ix = ctypes.c_void_p(_addr(idx) + 4 * int(row0[d]))
                self.xchg_gather_fixed(inp, width, ix, int(rows[d]), ctypes.c_void_p(outs[d]), stream)
        return 0

    def xchg_gather_bitmap(self, inp, bit_off, idx, n, out, null_count, stream):
        if n == 0:
            return 0
        ix = _arr(idx, np.uint32, n)
        bits = _bits(inp, bit_off, ix)
        packed = np.packbits(bits, bitorder="little")
        nbytes = ((n + 31) // 32) * 4
        o = _arr(out, np.uint8, nbytes)
        o[:] = 0
        o[: len(packed)] = packed
        if _addr(null_count):
            _arr(null_count, np.uint64, 1)[0] += int((~bits).sum())
        return 0

    def xchg_gather_offsets_workspace_bytes(self, n, out):
        out._obj.value = 64
        return 0

    def xchg_gather_offsets(self, in_off, idx, n, out_off, ws, wsb, stream):
        o = _arr(out_off, np.int64, n + 1)
        o[0] = 0
        if n == 0:
            return 0
        ix = _arr(idx, np.uint32, n).astype(np.int64)
        io = _arr(in_off, np.int64, int(ix.max()) + 2)
        o[1:] = np.cumsum(io[ix + 1] - io[ix])
        return 0

    def xchg_gather_bytes(self, in_data, in_off, idx, n, out_off, out_data, stream):
        if n == 0:
            return 0
        ix = _arr(idx, np.uint32, n).astype(np.int64)
        io = _arr(in_off, np.int64, int(ix.max()) + 2)
        oo = _arr(out_off, np.int64, n + 1)
        src = _arr(in_data, np.uint8, int(io[ix + 1].max()))
        dst = _arr(out_data, np.uint8, int(oo[n]))
        for j in range(n):
            dst[oo[j]:oo[j + 1]] = src[io[ix[j]]:io[ix[j] + 1]]
        return 0

    def xchg_gather_spans(self, in_off, idx, n, src_start, out_off, ws, wsb, stream):
        o = _arr(out_off, np.int64, n + 1)
        o[0] = 0
        if n == 0:
            return 0
        ix = _arr(idx, np.uint32, n).astype(np.int64)
        io = _arr(in_off, np.int64, int(ix.max()) + 2)
        _arr(src_start, np.int64, n)[:] = io[ix]
        o[1:] = np.cumsum(io[ix + 1] - io[ix])
        return 0

    def xchg_gather_bytes_spans(self, in_data, src_start, n, out_off, out_data, stream):
        if n == 0:
            return 0
        st = _arr(src_start, np.int64, n)
        oo = _arr(out_off, np.int64, n + 1)
        ln = np.diff(oo)
        src = _arr(in_data, np.uint8, int((st + ln).max()))
        dst = _arr(out_data, np.uint8, int(oo[n]))
        for j in range(n):
            dst[oo[j]:oo[j + 1]] = src[st[j]:st[j] + ln[j]]
        return 0

    def xchg_offsets_widen(self, inp, count, base, out, stream):
        i = _arr(inp, np.int32, count).astype(np.int64)
        _arr(out, np.int64, count)[:] = i - i[0] + base
        return 0

    def xchg_offsets_narrow(self, inp, count, out, stream):
        i = _arr(inp, np.int64, count)
        _arr(out, np.int32, count)[:] = (i - i[0]).astype(np.int32)
        return 0

    def xchg_offsets_uniform32(self, count, width, out, stream):
        _arr(out, np.int32, count)[:] = np.arange(count, dtype=np.int64) * width
        return 0

    def xchg_offsets_uniform64(self, count, width, base, out, stream):
        _arr(out, np.int64, count)[:] = base + np.arange(count, dtype=np.int64) * width
        return 0

    def xchg_offsets_check_uniform(self, off, n_rows, width, flag, stream):
        o = _arr(off, np.int64, n_rows + 1)
        if (np.diff(o) != width).any():
            _arr(flag, np.uint32, 1)[0] |= 1
        return 0

    # -- multi-GPU exchange: control path ---------------------------------------
    @staticmethod
    def _splitmix64(x):
        m = (1 << 64) - 1
        x = (x + 0x9E3779B97F4A7C15) & m
        x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & m
        x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & m
        return x ^ (x >> 31)

    def xchg_sample_keys(self, keys, kt, desc, validity, vbit, starts, rows, nblocks, per_block, seed,
                         first_block, out, capacity, count_out, stream):
        ks = [min(int(per_block), int(rows[b])) for b in range(nblocks)]
        total = sum(ks)
        count_out._obj.value = total
        if total == 0:
            return 0
        if total > capacity:
            self.err = b"xchg_sample_keys: samples exceed the capacity"
            return -4
        pos = 0
        for b in range(nblocks):
            k, r = ks[b], int(rows[b])
            if k == 0:
                continue
            sd = self._splitmix64(int(seed) ^ self._splitmix64(int(first_block) + b))
            j = np.arange(k, dtype=np.uint64)
            loc = j if k == r else self._feistel(sd, r, j, False)
            ridx = (int(starts[b]) + loc.astype(np.int64)).astype(np.uint32)
            tmp = np.ascontiguousarray(ridx)
            _, nk, _ = self._norm_col(keys, kt, desc, validity, vbit, ctypes.c_void_p(tmp.ctypes.data), k)
            _arr(out, np.uint64, total)[pos:pos + k] = nk
            pos += k
        return 0

    def xchg_select_boundaries(self, gathered, nmsgs, msg_words, count_word, hdr_words, R, bounds, info, stream):
        g = _arr(gathered, np.uint64, nmsgs * msg_words).reshape(nmsgs, msg_words)
        allv = np.concatenate([g[i, hdr_words:hdr_words + int(g[i, count_word])] for i in range(nmsgs)])
        total = len(allv)
        stride = (total + 16383) // 16384 if total > 16384 else 1
        used = np.sort(allv[::stride], kind="stable")
        m = len(used)
        if m and R > 1:
            step = 1.0 / float(R)
            idx = [min(m - 1, int((i * step) * float(m - 1))) for i in range(1, R)]
            _arr(bounds, np.uint64, R - 1)[:] = used[idx]
        if _addr(info):
            _arr(info, np.uint64, 2)[:] = (total, m)
        return 0

    @staticmethod
    def _dest(nk, thr, strict):
        """number of thresholds t with t < k, or t == k and not strict (thresholds sorted)"""
        if len(thr) == 0:
            return np.zeros(len(nk), dtype=np.int64)
        lt = np.searchsorted(thr, nk, side="left")       # thresholds < k
        d = lt.copy()
        # thresholds equal to k follow; the leading non-strict ones count (the rule is a prefix)
        for i in range(len(nk)):
            j = int(lt[i])
            while j < len(thr) and thr[j] == nk[i] and not strict[j]:
                j += 1
            d[i] = j
        return d

    def _dest_rows(self, keys, kt, desc, n, bounds, pick, strict_p, nthr):
        _, nk, _ = self._norm_col(keys, kt, desc, None, 0, None, n)
        picks = [int(pick[i]) for i in range(nthr)]
        b = _arr(bounds, np.uint64, max(picks) + 1) if nthr else np.zeros(0, np.uint64)
        thr = b[picks] if nthr else np.zeros(0, np.uint64)
        strict = np.array([int(strict_p[i]) if strict_p else 0 for i in range(nthr)], dtype=np.uint8)
        return nk, self._dest(nk, thr, strict)

    def xchg_partition_hist(self, keys, kt, desc, n, bounds, pick, strict, nthr, counts, stream):
        _arr(counts, np.uint64, nthr + 1)[:] = 0
        if n == 0:
            return 0
        _, d = self._dest_rows(keys, kt, desc, n, bounds, pick, strict, nthr)
        _arr(counts, np.uint64, nthr + 1)[:] = np.bincount(d, minlength=nthr + 1).astype(np.uint64)
        return 0

    def xchg_partition_workspace_bytes(self, n, out):
        out._obj.value = 64
        return 0

    def xchg_partition_rows(self, keys, kt, desc, n, bounds, pick, strict, nthr, counts, keys_out, rows_out,
                            ws, wsb, stream):
        if n == 0:
            return 0
        nk, d = self._dest_rows(keys, kt, desc, n, bounds, pick, strict, nthr)
        order = np.argsort(d, kind="stable")
        _arr(keys_out, np.uint64, n)[:] = nk[order]
        _arr(rows_out, np.uint32, n)[:] = order.astype(np.uint32)
        return 0

    def xchg_copy_segments(self, nseg, src, dst, nbytes, stream):
        for i in range(nseg):
            b = int(nbytes[i])
            if b:
                _arr(ctypes.c_void_p(dst[i]), np.uint8, b)[:] = _arr(ctypes.c_void_p(src[i]), np.uint8, b)
        return 0

    def xchg_signal_flags(self, ptrs, n, value, stream):
        for i in range(n):
            _arr(ctypes.c_void_p(ptrs[i]), np.uint32, 1)[0] = value
        return 0

    def xchg_wait_flags(self, flags, n, value, timeout_ns, error, stream):
        import time

        f = _arr(flags, np.uint32, n)
        t0 = time.monotonic()
        while True:
            late = [i for i in range(n) if ((int(f[i]) - int(value)) & 0xFFFFFFFF) >= 0x80000000]
            if not late:
                return 0
            if (time.monotonic() - t0) * 1e9 > timeout_ns:
                _arr(error, np.uint32, 1)[0] = 1 + late[0]
                return 0
            time.sleep(0.0005)

    # -- merge ---------------------------------------------------------------------
    def xchg_merge_workspace_bytes(self, n, out):
        out._obj.value = n * 12 + 1024
        return 0

    def xchg_merge_sorted_runs(self, keys_in, vals_in, run_start, nruns, keys_out, vals_out, ws, wsb, flag, stream):
        """Contract: the stable sort of the concatenation (what merging sorted runs with ties
        resolved by (run, position) gives); the flag reports a run that is not ascending."""
        if _addr(flag):
            _arr(flag, np.uint32, 1)[0] = 0
        if nruns == 0:
            return 0
        n = int(run_start[nruns])
        if n == 0:
            return 0
        if nruns > 128:
            self.err = b"at most 128 runs per call"
            return -1
        k = _arr(keys_in, np.uint64, n)
        v = _arr(vals_in, np.uint32, n) if _addr(vals_in) else np.arange(n, dtype=np.uint32)
        if _addr(flag):
            bad = False
            for r in range(nruns):
                seg = k[int(run_start[r]):int(run_start[r + 1])]
                bad = bad or bool((seg[1:] < seg[:-1]).any())
            _arr(flag, np.uint32, 1)[0] = int(bad)
        order = np.argsort(k, kind="stable")
        _arr(keys_out, np.uint64, n)[:] = k[order]
        _arr(vals_out, np.uint32, n)[:] = v[order]
        return 0

    # -- exchange-side entry points ------------------------------------------
    def xchg_assign_partitions(self, cols, col_desc, ncols, n, lower, upper, flags, nb, first_desc, dest, counts,
                               stream):
        """Independent statement: sort the rows (stable, lexicographic over the
        normalised key columns), cut the sorted order exactly as
        xchg_partition_bounds does, and give every row the partition its sorted
        position falls into."""
        if _addr(counts):
            _arr(counts, np.uint64, nb + 1)[:] = 0
        if n == 0:
            return 0
        arrays = [_arr(cols[c], np.uint64, n) for c in range(ncols)]
        order = np.lexsort(tuple(arrays[::-1]))  # stable; first column is the primary key
        sorted_cols = [a[order] for a in arrays]
        lo_a = _arr(lower, np.uint64, nb * ncols)
        up_a = _arr(upper, np.uint64, nb * ncols)
        fl_a = _arr(flags, np.uint8, nb * ncols) if _addr(flags) else np.zeros(nb * ncols, np.uint8)
        cuts = []
        for b in range(nb):
            lo, hi = 0, n
            early = False
            for c in range(ncols):
                if lo == hi:
                    early = True
                    break
                f = int(fl_a[b * ncols + c])
                if f & 2:
                    break
                if f & 4:
                    if col_desc[c]:
                        hi = lo
                    else:
                        lo = hi
                    continue
                seg = sorted_cols[c][lo:hi]
                l2 = lo + int(np.searchsorted(seg, lo_a[b * ncols + c], side="left"))
                h2 = l2 if (f & 1) else lo + int(np.searchsorted(seg, up_a[b * ncols + c], side="right"))
                lo, hi = l2, h2
            cuts.append(hi if (early or first_desc) else lo)
        pos_part = np.searchsorted(np.asarray(cuts, dtype=np.int64), np.arange(n), side="right") if nb else np.zeros(n, np.int64)
        d = np.empty(n, dtype=np.uint32)
        d[order] = pos_part.astype(np.uint32)
        _arr(dest, np.uint32, n)[:] = d
        if _addr(counts):
            _arr(counts, np.uint64, nb + 1)[:] = np.bincount(d, minlength=nb + 1).astype(np.uint64)
        return 0

    def xchg_bitmap_unpack(self, inp, bit_off, n, out, stream):
        if n == 0:
            return 0
        _arr(out, np.uint8, n)[:] = _bits(inp, bit_off, np.arange(n, dtype=np.int64)).astype(np.uint8)
        return 0

    def xchg_bitmap_pack(self, inp, n, out, zero_count, stream):
        if n == 0:
            return 0
        bits = _arr(inp, np.uint8, n) != 0
        packed = np.packbits(bits, bitorder="little")
        o = _arr(out, np.uint8, ((n + 31) // 32) * 4)
        o[:] = 0
        o[: len(packed)] = packed
        if _addr(zero_count):
            _arr(zero_count, np.uint64, 1)[0] += int((~bits).sum())
        return 0

    def xchg_offsets_rebase(self, inp, count, base, out, stream):
        if count == 0:
            return 0
        i = _arr(inp, np.int64, count)
        _arr(out, np.int64, count)[:] = i - i[0] + base
        return 0

    # -- shuffle -----------------------------------------------------------------
    @staticmethod
    def _philox2x32_10(c0, c1, key):
        c0 = c0.astype(np.uint64)
        c1 = c1.astype(np.uint64)
        key = np.uint64(key)
        M = np.uint64(0xD256D193)
        m32 = np.uint64(0xFFFFFFFF)
        for _ in range(10):
            p = M * c0
            hi, lo = p >> np.uint64(32), p & m32
            c0 = (hi ^ key ^ c1) & m32
            c1 = lo
            key = (key + np.uint64(0x9E3779B9)) & m32
        return c0

    def _feistel(self, seed, n_total, x, inverse):
        bits = 2
        while (1 << bits) < n_total:
            bits += 1
        bits += bits & 1
        hb = bits // 2
        mask = np.uint64((1 << hb) - 1)
        key, tweak = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
        x = x.astype(np.uint64).copy()
        todo = np.ones(len(x), dtype=bool)
        while todo.any():
            v = x[todo]
            L, R = v >> np.uint64(hb), v & mask
            rounds = range(4) if not inverse else range(3, -1, -1)
            for r in rounds:
                tw = np.full(len(v), (tweak ^ ((0x9E3779B9 * (r + 1)) & 0xFFFFFFFF)), dtype=np.uint64)
                if not inverse:
                    t = L ^ (self._philox2x32_10(R, tw, key) & mask)
                    L, R = R, t
                else:
                    t = R ^ (self._philox2x32_10(L, tw, key) & mask)
                    R, L = L, t
            v = (L << np.uint64(hb)) | R
            x[todo] = v
            todo[todo] = v >= np.uint64(n_total)
        return x

    def xchg_shuffle_positions(self, seed, n_total, first, count, inverse, out, stream):
        if count == 0:
            return 0
        x = np.arange(first, first + count, dtype=np.uint64)
        _arr(out, np.uint32, count)[:] = self._feistel(int(seed), int(n_total), x, bool(inverse)).astype(np.uint32)
        return 0

    def xchg_invert_permutation(self, pos, n, base, inv, stream):
        if n == 0:
            return 0
        p = _arr(pos, np.uint32, n).astype(np.int64) - base
        _arr(inv, np.uint32, n)[p] = np.arange(n, dtype=np.uint32)
        return 0

    def xchg_even_split_dest(self, pos, n, n_total, parts, dest, counts, stream):
        _arr(counts, np.uint64, parts)[:] = 0
        if n == 0:
            return 0
        p = _arr(pos, np.uint32, n).astype(np.int64)
        starts = np.array([(k * n_total) // parts for k in range(parts)], dtype=np.int64)
        d = np.searchsorted(starts, p, side="right") - 1
        _arr(dest, np.uint32, n)[:] = d.astype(np.uint32)
        _arr(counts, np.uint64, parts)[:] = np.bincount(d, minlength=parts).astype(np.uint64)
        return 0

    # -- aggregate ---------------------------------------------------------------
    # Host-memory statement of the aggregate entry points' CONTRACT (numpy; no hashing:
    # "slot" is simply the index of the key among the distinct keys, shuffled so that
    # callers cannot rely on table order).
    _EMPTY = np.uint64(0xFFFFFFFFFFFFFFFF)

    def xchg_hash_table_bytes(self, capacity, out):
        out._obj.value = capacity * 8 + (capacity + 2) * 4 + 1024
        return 0

    def xchg_hash_groups(self, norm, validity, vbit, n, capacity, ws, ws_bytes, slot_out, gkeys, gslots,
                         ngroups, stream):
        keys = _arr(norm, np.uint64, n)
        valid = _bits(validity, vbit, np.arange(n, dtype=np.int64)) if _addr(validity) and n else np.ones(n, bool)
        regular = valid & (keys != self._EMPTY)
        uniq = np.unique(keys[regular])
        if len(uniq) > capacity:
            self.err = b"hash table overflowed"
            return -4
        rng = np.random.default_rng(len(uniq))
        slots_of_uniq = rng.permutation(capacity)[: len(uniq)].astype(np.uint32)
        slot = np.empty(n, dtype=np.uint32)
        slot[regular] = slots_of_uniq[np.searchsorted(uniq, keys[regular])]
        slot[~valid] = capacity
        is_max = valid & (keys == self._EMPTY)
        slot[is_max] = capacity + 1
        _arr(slot_out, np.uint32, n)[:] = slot
        order = np.argsort(slots_of_uniq, kind="stable")  # table order
        gk = list(uniq[order])
        gs = list(slots_of_uniq[order])
        if is_max.any():
            gk.append(self._EMPTY)
            gs.append(capacity + 1)
        if (~valid).any():
            gk.append(self._EMPTY)
            gs.append(capacity)
        g = len(gk)
        _arr(gkeys, np.uint64, g)[:] = np.array(gk, dtype=np.uint64)
        _arr(gslots, np.uint32, g)[:] = np.array(gs, dtype=np.uint32)
        ngroups._obj.value = g
        self._gid_of_slot = None
        return 0

    def xchg_group_ids(self, order, gslots, g, slot, n, capacity, ws, rgid, stream):
        if g == 0 or n == 0:
            return 0
        o = _arr(order, np.uint32, g).astype(np.int64)
        s = _arr(gslots, np.uint32, g).astype(np.int64)
        gid_of_slot = np.full(capacity + 2, 0xFFFFFFFF, dtype=np.uint32)
        gid_of_slot[s[o]] = np.arange(g, dtype=np.uint32)
        _arr(rgid, np.uint32, n)[:] = gid_of_slot[_arr(slot, np.uint32, n)]
        return 0

    def xchg_agg_stats_init(self, stats, m2, g, stream):
        if g == 0:
            return 0
        st = _arr(stats, np.uint64, 5 * g)
        st[:] = 0
        st[2 * g:3 * g] = self._EMPTY
        if _addr(m2):
            _arr(m2, np.float64, g)[:] = 0
        return 0

    @staticmethod
    def _gids(rgid, n):
        return _arr(rgid, np.uint32, n).astype(np.int64) if _addr(rgid) else np.zeros(n, np.int64)

    def xchg_key_minmax(self, norm, validity, vbit, n, scratch, out3, stream):
        out3[0], out3[1], out3[2] = int(self._EMPTY), 0, 0
        if n == 0:
            return 0
        keys = _arr(norm, np.uint64, n)
        valid = _bits(validity, vbit, np.arange(n, dtype=np.int64)) if _addr(validity) else np.ones(n, bool)
        if valid.any():
            out3[0], out3[1], out3[2] = int(keys[valid].min()), int(keys[valid].max()), int(valid.sum())
        return 0

    def xchg_dense_groups_bytes(self, rng, out):
        if rng == 0 or rng > (1 << 28):
            self.err = b"dense key range out of bounds"
            return -3
        out._obj.value = rng * 5 + 4096
        return 0

    def xchg_dense_groups(self, norm, validity, vbit, n, kmin, rng, ws, wsb, rgid, gkeys, grows, ngroups, stream):
        ngroups._obj.value = 0
        if n == 0:
            return 0
        keys = _arr(norm, np.uint64, n)
        valid = _bits(validity, vbit, np.arange(n, dtype=np.int64)) if _addr(validity) else np.ones(n, bool)
        uniq = np.unique(keys[valid])
        assert len(uniq) == 0 or (int(uniq[0]) >= kmin and int(uniq[-1]) < kmin + rng)
        gid = np.full(n, len(uniq), dtype=np.uint32)
        gid[valid] = np.searchsorted(uniq, keys[valid]).astype(np.uint32)
        _arr(rgid, np.uint32, n)[:] = gid
        g = len(uniq)
        _arr(gkeys, np.uint64, g)[:] = uniq
        if (~valid).any():
            _arr(gkeys, np.uint64, g + 1)[g] = self._EMPTY
            g += 1
        if _addr(grows):
            _arr(grows, np.uint64, g)[:] = np.bincount(gid.astype(np.int64), minlength=g).astype(np.uint64)
        ngroups._obj.value = g
        return 0

    def xchg_group_flags_fixed(self, norm, validity, vbit, perm, n, flags, stream):
        if n == 0:
            return 0
        k = _arr(norm, np.uint64, n)
        fl = _arr(flags, np.uint8, n)
        diff = np.ones(n, bool)
        diff[1:] = k[1:] != k[:-1]
        if _addr(validity):
            p = _arr(perm, np.uint32, n).astype(np.int64) if _addr(perm) else np.arange(n, dtype=np.int64)
            v = _bits(validity, vbit, p)
            diff[1:] |= v[1:] != v[:-1]
        fl[diff] = 1
        return 0

    def xchg_group_flags_bytes(self, off, data, validity, vbit, perm, n, flags, stream):
        if n == 0:
            return 0
        p = _arr(perm, np.uint32, n).astype(np.int64) if _addr(perm) else np.arange(n, dtype=np.int64)
        nrows = int(p.max()) + 1
        rows = self._rows_bytes(off, data, nrows)
        v = _bits(validity, vbit, p) if _addr(validity) else np.ones(n, bool)
        fl = _arr(flags, np.uint8, n)
        fl[0] = 1
        for i in range(1, n):
            a, b = int(p[i - 1]), int(p[i])
            if v[i] != v[i - 1] or (v[i] and rows[a] != rows[b]):
                fl[i] = 1
        return 0

    def xchg_group_first_rows(self, rgid, n, g, first, stream):
        if g == 0:
            return 0
        out = _arr(first, np.uint32, g)
        out[:] = 0xFFFFFFFF
        if n:
            np.minimum.at(out, _arr(rgid, np.uint32, n).astype(np.int64), np.arange(n, dtype=np.uint32))
        return 0

    def xchg_groups_from_flags_bytes(self, n, out):
        out._obj.value = (n // 4096 + 2) * 4 + 1024
        return 0

    def xchg_groups_from_flags(self, flags, perm, n, ws, wsb, rgid, first, ngroups, stream):
        ngroups._obj.value = 0
        if n == 0:
            return 0
        fl = _arr(flags, np.uint8, n).astype(np.int64)
        gid = np.cumsum(fl) - 1
        p = _arr(perm, np.uint32, n).astype(np.int64) if _addr(perm) else np.arange(n, dtype=np.int64)
        _arr(rgid, np.uint32, n)[p] = gid.astype(np.uint32)
        g = int(gid[-1]) + 1
        _arr(first, np.uint32, g)[:] = p[fl == 1].astype(np.uint32)
        ngroups._obj.value = g
        return 0

    def xchg_agg_count_rows(self, rgid, n, num_groups, rows, stream):
        if n == 0:
            return 0
        g = self._gids(rgid, n)
        r = _arr(rows, np.uint64, int(g.max()) + 1)
        r += np.bincount(g, minlength=len(r)).astype(np.uint64)
        return 0

    @staticmethod
    def _enc(x, dt):
        if dt.kind == "f":
            b = x.astype(np.float64).view(np.uint64)
            return np.where(b >> np.uint64(63), ~b, b | np.uint64(1 << 63))
        if dt.kind == "i":
            return x.astype(np.int64).view(np.uint64) ^ np.uint64(1 << 63)
        return x.astype(np.uint64)

    def xchg_agg_accumulate(self, rgid, values, vt, validity, vbit, n, g, stats, want_minmax, rows, stream):
        if n == 0:
            return 0
        dt = np.dtype(_KT[vt])
        gid = self._gids(rgid, n)
        x = _arr(values, dt, n)
        valid = _bits(validity, vbit, np.arange(n, dtype=np.int64)) if _addr(validity) else np.ones(n, bool)
        st = _arr(stats, np.uint64, 5 * g)
        cnt, mn, mx, nan = st[0:g], st[2 * g:3 * g], st[3 * g:4 * g], st[4 * g:5 * g]
        cnt += np.bincount(gid[valid], minlength=g).astype(np.uint64)
        if dt.kind == "f":
            s = st[g:2 * g].view(np.float64)
            with np.errstate(invalid="ignore", over="ignore"):
                np.add.at(s, gid[valid], x[valid].astype(np.float64))
        else:
            s = st[g:2 * g].view(np.int64)
            np.add.at(s, gid[valid], x[valid].astype(np.int64) if dt != np.uint64 else x[valid].view(np.int64))
        if want_minmax:
            isnan = np.isnan(x) if dt.kind == "f" else np.zeros(n, bool)
            nan += np.bincount(gid[valid & isnan], minlength=g).astype(np.uint64)
            ok = valid & ~isnan
            e = self._enc(x[ok], dt)
            np.minimum.at(mn, gid[ok], e)
            np.maximum.at(mx, gid[ok], e)
        return 0

    def xchg_agg_accumulate_m2(self, rgid, values, vt, validity, vbit, n, g, stats, m2, stream):
        if n == 0:
            return 0
        dt = np.dtype(_KT[vt])
        gid = self._gids(rgid, n)
        x = _arr(values, dt, n).astype(np.float64)
        valid = _bits(validity, vbit, np.arange(n, dtype=np.int64)) if _addr(validity) else np.ones(n, bool)
        st = _arr(stats, np.uint64, 5 * g)
        cnt = st[0:g].astype(np.float64)
        s = st[g:2 * g].view(np.float64) if dt.kind == "f" else st[g:2 * g].view(np.int64).astype(np.float64)
        with np.errstate(invalid="ignore", divide="ignore"):
            mean = s / cnt
            d = x[valid] - mean[gid[valid]]
            np.add.at(_arr(m2, np.float64, g), gid[valid], d * d)
        return 0

    def xchg_agg_merge_rows(self, rgid, part_rows, n, rows, stream):
        if n == 0:
            return 0
        g = self._gids(rgid, n)
        r = _arr(rows, np.uint64, int(g.max()) + 1)
        np.add.at(r, g, _arr(part_rows, np.int64, n).astype(np.uint64))
        return 0

    def xchg_agg_merge_states(self, rgid, pst, n, sum_is_int, g, stats, m2, stream):
        if n == 0:
            return 0
        gid = self._gids(rgid, n)
        p = _arr(pst, np.uint64, 6 * n).reshape(n, 6)
        st = _arr(stats, np.uint64, 5 * g)
        has = p[:, 0] > 0
        np.add.at(st[0:g], gid[has], p[has, 0])
        if sum_is_int:
            np.add.at(st[g:2 * g].view(np.int64), gid[has], p[has, 1].view(np.int64))
        else:
            with np.errstate(invalid="ignore", over="ignore"):
                np.add.at(st[g:2 * g].view(np.float64), gid[has], p[has, 1].view(np.float64))
        np.minimum.at(st[2 * g:3 * g], gid[has], p[has, 2])
        np.maximum.at(st[3 * g:4 * g], gid[has], p[has, 3])
        np.add.at(st[4 * g:5 * g], gid[has], p[has, 4])
        if _addr(m2):
            cnt = st[0:g].astype(np.float64)
            tot = st[g:2 * g].view(np.int64).astype(np.float64) if sum_is_int else st[g:2 * g].view(np.float64)
            si = p[:, 1].view(np.int64).astype(np.float64) if sum_is_int else p[:, 1].view(np.float64)
            with np.errstate(invalid="ignore", divide="ignore"):
                d = si[has] / p[has, 0].astype(np.float64) - (tot / cnt)[gid[has]]
                np.add.at(_arr(m2, np.float64, g), gid[has],
                          p[has, 5].view(np.float64) + p[has, 0].astype(np.float64) * d * d)
        return 0

    def xchg_agg_pack_states(self, g, stats, m2, out, stream):
        if g == 0:
            return 0
        st = _arr(stats, np.uint64, 5 * g).reshape(5, g)
        o = _arr(out, np.uint64, 6 * g).reshape(g, 6)
        o[:, :5] = st.T
        o[:, 5] = _arr(m2, np.float64, g).view(np.uint64) if _addr(m2) else 0
        return 0

    def xchg_agg_finalize(self, kind, ddof, ignore_nulls, vt, g, rows, stats, m2, out, valid_out, stream):
        if g == 0:
            return 0
        if not 0 <= kind <= 6:
            self.err = b"kind must be 0..6"
            return -1
        r = _arr(rows, np.uint64, g)
        o = _arr(out, np.uint64, g)
        v = _arr(valid_out, np.uint8, g)
        if kind == 0:
            o[:] = r
            v[:] = 1
            return 0
        dt = np.dtype(_KT[vt])
        is_int = dt.kind in "iu"
        st = _arr(stats, np.uint64, 5 * g).reshape(5, g)
        cnt = st[0]
        if kind == 6 and not is_int:
            cnt = cnt - np.minimum(st[4], cnt)  # AbsMax: NaN counts as null (pandas.isnull)
        ok = (cnt > 0) & (bool(ignore_nulls) | (cnt == r))
        v[:] = ok.astype(np.uint8)
        s = st[1].view(np.int64).astype(np.float64) if is_int else st[1].view(np.float64)
        res = np.zeros(g, dtype=np.uint64)
        with np.errstate(invalid="ignore", divide="ignore"):
            if kind == 1:
                res = st[1].copy()
            elif kind in (2, 3):
                e = st[2] if kind == 2 else st[3]
                if is_int:
                    res = e ^ np.uint64(1 << 63) if dt.kind == "i" else e.copy()
                else:
                    b = np.where(e >> np.uint64(63), e & np.uint64((1 << 63) - 1), ~e)
                    f = b.view(np.float64).copy()
                    f[st[4] >= cnt] = np.nan
                    res = f.view(np.uint64)
            elif kind == 4:
                res = (s / cnt.astype(np.float64)).view(np.uint64)
            elif kind == 6:
                def dec(e):
                    b = np.where(e >> np.uint64(63), e & np.uint64((1 << 63) - 1), ~e)
                    return b.view(np.float64)

                if is_int and dt.kind == "i":
                    mn = (st[2] ^ np.uint64(1 << 63)).view(np.int64).astype(object)
                    mx = (st[3] ^ np.uint64(1 << 63)).view(np.int64).astype(object)
                    res = np.array([max(abs(int(x)), abs(int(y))) & ((1 << 64) - 1) for x, y in zip(mn, mx)],
                                   dtype=np.uint64)
                elif is_int:
                    res = st[3].copy()
                else:
                    f = np.maximum(np.abs(dec(st[2])), np.abs(dec(st[3])))
                    res = np.ascontiguousarray(f).view(np.uint64)
            else:
                m = _arr(m2, np.float64, g)
                c = cnt.astype(np.float64)
                val = np.where(c >= 2, np.sqrt(m / (c - ddof)), 0.0)
                res = val.view(np.uint64)
        o[:] = np.where(ok, res, np.uint64(0))
        return 0

    def xchg_group_keys_out(self, norm, kt, g, out, stream):
        if g == 0:
            return 0
        k = _arr(norm, np.uint64, g)
        dt = np.dtype(_KT[kt])
        o = _arr(out, np.uint64, g)
        bits = dt.itemsize * 8
        if dt.kind == "i":
            u = k & np.uint64((1 << bits) - 1) if bits < 64 else k
            x = (u ^ np.uint64(1 << (bits - 1))).astype(np.uint64)
            if bits < 64:
                x = x.astype(np.int64)
                x = np.where(x >= (1 << (bits - 1)), x - (1 << bits), x)
                o[:] = x.astype(np.int64).view(np.uint64)
            else:
                o[:] = x
        elif dt.kind == "u":
            o[:] = k
        else:
            nan_key = np.uint64((1 << bits) - 2) if bits < 64 else np.uint64(0xFFFFFFFFFFFFFFFE)
            if bits == 32:
                t = (k & np.uint64(0xFFFFFFFF)).astype(np.uint32)
                b = np.where(t & np.uint32(0x80000000), t & np.uint32(0x7FFFFFFF), ~t)
                f = b.view(np.float32).astype(np.float64)
            else:
                b = np.where(k >> np.uint64(63), k & np.uint64((1 << 63) - 1), ~k)
                f = b.view(np.float64).copy()
            f = np.where(k == nan_key, np.nan, f)
            o[:] = f.view(np.uint64)
        return 0

    def xchg_gather_bitmap_bytes(self, inp, bit_off, idx, n, out, stream):
        if n == 0:
            return 0
        _arr(out, np.uint8, n)[:] = _bits(inp, bit_off, _arr(idx, np.uint32, n)).astype(np.uint8)
        return 0

    def xchg_set_gather_stage_kb(self, kb):
        return 0

    def xchg_set_gather_ctas_per_sm(self, n):
        return 0

    def xchg_set_sort_algorithm(self, a):
        return 0 if a in (0, 1) else 1

    # -- string keys -------------------------------------------------------------
    @staticmethod
    def _rows_bytes(off, data, n):
        o = _arr(off, np.int64, n + 1)
        d = _arr(data, np.uint8, int(o[n])) if n and int(o[n]) else np.zeros(0, np.uint8)
        return [bytes(d[o[i]:o[i + 1]]) for i in range(n)]

    def xchg_string_lengths(self, off, n, len_out, max_len, stream):
        _arr(max_len, np.uint64, 1)[0] = 0
        if n == 0:
            return 0
        o = _arr(off, np.int64, n + 1)
        l = np.diff(o)
        _arr(len_out, np.uint32, n)[:] = np.minimum(l, 0xFFFFFFFF).astype(np.uint32)
        _arr(max_len, np.uint64, 1)[0] = int(l.max())
        return 0

    def xchg_string_chunk_keys(self, off, data, n, chunk, out, stream):
        if n == 0:
            return 0
        rows = self._rows_bytes(off, data, n)
        _arr(out, np.uint64, n)[:] = np.array(
            [int.from_bytes(r[8 * chunk: 8 * chunk + 8].ljust(8, b"\0"), "big") for r in rows], dtype=np.uint64)
        return 0

    def _str_inputs(self, off, data, validity, n, boff, bdata, bflags, nb):
        rows = self._rows_bytes(off, data, n)
        valid = _bits(validity, 0, np.arange(n, dtype=np.int64)) if _addr(validity) and n else np.ones(n, bool)
        bounds = self._rows_bytes(boff, bdata, nb)
        fl = _arr(bflags, np.uint8, nb) if _addr(bflags) else np.zeros(nb, np.uint8)
        return rows, valid, bounds, fl

    # -- MSD refinement of a string sort --------------------------------------------
    def _row_bytes(self, off, data, r):
        o = _arr(off, np.int64, r + 2)
        return bytes(_arr(data, np.uint8, int(o[r + 1]))[int(o[r]):int(o[r + 1])]) if o[r + 1] > o[r] else b""

    def xchg_string_tie_flags(self, off, data, validity, rows, keys, gid_prev, m, from_byte, gflag, fflag, stream):
        if m == 0:
            return 0
        r = _arr(rows, np.uint32, m).astype(np.int64)
        k = _arr(keys, np.uint64, m)
        gp = _arr(gid_prev, np.uint32, m) if _addr(gid_prev) else None
        valid = _bits(validity, 0, r) if _addr(validity) else np.ones(m, bool)
        g = np.zeros(m, np.uint8)
        f = np.zeros(m, np.uint8)
        g[0] = f[0] = 1
        for j in range(1, m):
            gj = (gp is not None and gp[j] != gp[j - 1]) or k[j] != k[j - 1] or valid[j] != valid[j - 1]
            fj = gj
            if not gj and valid[j]:
                a, b = self._row_bytes(off, data, int(r[j - 1])), self._row_bytes(off, data, int(r[j]))
                fj = len(a) != len(b) or a[from_byte:] != b[from_byte:]
            g[j], f[j] = int(gj), int(fj)
        _arr(gflag, np.uint8, m)[:] = g
        _arr(fflag, np.uint8, m)[:] = f
        return 0

    def xchg_string_active_rows(self, gflag, fflag, gid, m, ngroups, scratch, active, stream):
        if m == 0:
            return 0
        g, f = _arr(gflag, np.uint8, m), _arr(fflag, np.uint8, m)
        gi = _arr(gid, np.uint32, m).astype(np.int64)
        unresolved = np.zeros(ngroups, bool)
        unresolved[gi[(f != 0) & (g == 0)]] = True
        _arr(active, np.uint32, m)[:] = unresolved[gi].astype(np.uint32)
        return 0

    def xchg_string_run_heads(self, fflag, active, m, period, head, stream):
        if m == 0:
            return 0
        f = _arr(fflag, np.uint8, m) != 0
        if _addr(active):
            f = f | (_arr(active, np.uint32, m) != 0)
        if period:
            f = f | (np.arange(m) % period == 0)
        _arr(head, np.uint8, m)[:] = f.astype(np.uint8)
        return 0

    def xchg_scatter_flagged(self, flags, excl, m, out, stream):
        if m == 0:
            return 0
        fl = _arr(flags, np.uint32, m)
        idx = np.nonzero(fl)[0].astype(np.uint32)
        _arr(out, np.uint32, len(idx))[:] = idx
        return 0

    def xchg_string_chunk_keys_rows(self, off, data, rows, m, chunk, out, stream):
        if m == 0:
            return 0
        r = _arr(rows, np.uint32, m).astype(np.int64)
        res = np.zeros(m, np.uint64)
        for j in range(m):
            b = self._row_bytes(off, data, int(r[j]))
            if chunk == 0xFFFFFFFF:
                res[j] = len(b)
            else:
                res[j] = int.from_bytes(b[8 * chunk:8 * chunk + 8].ljust(8, b"\0"), "big")
        _arr(out, np.uint64, m)[:] = res
        return 0

    def xchg_scatter_u32(self, idx, vals, m, out, stream):
        if m == 0:
            return 0
        ix = _arr(idx, np.uint32, m).astype(np.int64)
        _arr(out, np.uint32, int(ix.max()) + 1)[ix] = _arr(vals, np.uint32, m)
        return 0

    def xchg_partition_bounds_bytes(self, off, data, validity, n, boff, bdata, bflags, nb, desc, cuts, stream):
        rows, valid, bounds, fl = self._str_inputs(off, data, validity, n, boff, bdata, bflags, nb)
        out = _arr(cuts, np.uint64, nb)
        for b in range(nb):
            none = bool(fl[b] & 4)
            if not desc:
                lo, hi = 0, n
                while lo < hi:
                    mid = (lo + hi) // 2
                    before = False if not valid[mid] else (True if none else rows[mid] < bounds[b])
                    lo, hi = (mid + 1, hi) if before else (lo, mid)
                out[b] = lo
            else:
                mn, mx = 0, n
                while mn < mx:
                    mid = mn + (mx - mn) // 2
                    r = n - 1 - mid
                    el = not valid[r]
                    lt = (not el) if none else ((not el) and rows[r] < bounds[b])
                    mn, mx = (mid + 1, mx) if lt else (mn, mid)
                out[b] = n - mn
        return 0

    def xchg_host_register(self, ptr, nbytes, already_out):
        return 0  # host memory is host memory here

    def xchg_host_unregister(self, ptr):
        return 0

    # -- key tuples with string columns ------------------------------------------------
    class _Last:
        """NULL_SENTINEL of the reference (util.py:67-89): greater than every value, equal
        to itself only."""

        def __eq__(self, o):
            return isinstance(o, type(self))

        def __lt__(self, o):
            return False

        def __le__(self, o):
            return isinstance(o, type(self))

        def __gt__(self, o):
            return not self.__le__(o)

        def __ge__(self, o):
            return True

        __hash__ = object.__hash__

    def _mixed_inputs(self, cols, ncols, n, nb):
        out = []
        for c in range(ncols):
            d = cols[c]
            if d.kind == 1:
                rows = self._rows_bytes(d.offsets, d.data, n)
                valid = (_bits(d.validity, 0, np.arange(n, dtype=np.int64)) if d.validity and n
                         else np.ones(n, bool))
                out.append(("s", bool(d.descending), rows, valid, self._rows_bytes(d.bound_offsets, d.bound_data, nb)))
            else:
                out.append(("n", bool(d.descending), _arr(d.norm, np.uint64, n), int(d.last_threshold),
                            bool(d.has_last)))
        return out

    def xchg_partition_bounds_mixed(self, cols, ncols, n, lower, upper, flags, nb, first_desc, cuts, stream):
        """Independent statement: string columns as the reference sees them -- object arrays
        with NULL_SENTINEL, numpy.searchsorted with the reversing sorter for descending
        columns (util.py:761-802); numeric columns through xchg_partition_bounds' own
        narrowing, one column at a time."""
        if nb == 0:
            return 0
        info = self._mixed_inputs(cols, ncols, n, nb)
        fl_a = _arr(flags, np.uint8, nb * ncols) if _addr(flags) else np.zeros(nb * ncols, np.uint8)
        out = _arr(cuts, np.uint64, nb)
        last = self._Last()
        for b in range(nb):
            lo, hi, early = 0, n, False
            for c in range(ncols):
                if lo == hi:
                    early = True
                    break
                f = int(fl_a[b * ncols + c])
                if f & 2:
                    break
                if info[c][0] == "s":
                    _, desc, rows, valid, bounds = info[c]
                    vals = np.empty(hi - lo, dtype=object)
                    for j in range(lo, hi):
                        vals[j - lo] = rows[j] if valid[j] else last
                    key = last if (f & 4) else bounds[b]
                    m = hi - lo
                    if desc:
                        srt = np.arange(m - 1, -1, -1)
                        nl = lo + (m - int(np.searchsorted(vals, key, side="right", sorter=srt)))
                        nh = lo + (m - int(np.searchsorted(vals, key, side="left", sorter=srt)))
                    else:
                        nl = lo + int(np.searchsorted(vals, key, side="left"))
                        nh = lo + int(np.searchsorted(vals, key, side="right"))
                    lo, hi = nl, nh
                    continue
                # one numeric column: run the numeric entry point on the run [lo, hi)
                _, desc, norm, thr, has_last = info[c]
                seg = np.ascontiguousarray(norm[lo:hi])
                one = (ctypes.c_void_p * 1)(seg.ctypes.data)
                cd = (ctypes.c_uint8 * 1)(int(desc))
                th = (ctypes.c_uint64 * 1)(thr)
                hl = (ctypes.c_uint8 * 1)(int(has_last))
                e = b * ncols + c
                L = np.array([_arr(lower, np.uint64, nb * ncols)[e]], dtype=np.uint64)
                U = np.array([_arr(upper, np.uint64, nb * ncols)[e]], dtype=np.uint64)
                F = np.array([f], dtype=np.uint8)
                res = np.zeros(2, dtype=np.uint64)
                # first_desc=0 -> left edge, first_desc=1 -> right edge of the equal range
                for side in (0, 1):
                    self.xchg_partition_bounds(one, cd, th, hl, 1, hi - lo, L.ctypes.data, U.ctypes.data,
                                               F.ctypes.data, 1, side, res[side:].ctypes.data, None)
                lo, hi = lo + int(res[0]), lo + int(res[1])
            out[b] = hi if (early or first_desc) else lo
        return 0

    def xchg_assign_partitions_mixed(self, cols, ncols, n, lower, upper, flags, nb, first_desc, dest, counts,
                                     stream):
        """Per row: count the boundaries the row is right of (tuple compare in sort order)."""
        if _addr(counts):
            _arr(counts, np.uint64, nb + 1)[:] = 0
        if n == 0:
            return 0
        info = self._mixed_inputs(cols, ncols, n, nb)
        lo_a = _arr(lower, np.uint64, nb * ncols)
        up_a = _arr(upper, np.uint64, nb * ncols)
        fl_a = _arr(flags, np.uint8, nb * ncols) if _addr(flags) else np.zeros(nb * ncols, np.uint8)

        def cmp(i, b):
            for c in range(ncols):
                f = int(fl_a[b * ncols + c])
                if f & 2:
                    return 0
                if info[c][0] == "s":
                    _, desc, rows, valid, bounds = info[c]
                    none = bool(f & 4)
                    if not desc:
                        if not valid[i]:
                            if none:
                                continue
                            return 1
                        if none:
                            return -1
                    elif none or not valid[i]:
                        return 1
                    r = (rows[i] > bounds[b]) - (rows[i] < bounds[b])
                    if r:
                        return -r if desc else r
                    continue
                _, desc, norm, _, _ = info[c]
                if f & 4:
                    return 1 if desc else -1
                t = int(norm[i])
                if t < int(lo_a[b * ncols + c]):
                    return -1
                if (f & 1) or t > int(up_a[b * ncols + c]):
                    return 1
            return 0

        d = np.zeros(n, dtype=np.uint32)
        for i in range(n):
            cnt = 0
            for b in range(nb):
                c = cmp(i, b)
                if (c > 0) if first_desc else (c >= 0):
                    cnt += 1
                else:
                    break
            d[i] = cnt
        _arr(dest, np.uint32, n)[:] = d
        if _addr(counts):
            _arr(counts, np.uint64, nb + 1)[:] = np.bincount(d, minlength=nb + 1).astype(np.uint64)
        return 0

    def xchg_assign_partitions_bytes(self, off, data, validity, n, boff, bdata, bflags, nb, desc, dest, counts,
                                     stream):
        if _addr(counts):
            _arr(counts, np.uint64, nb + 1)[:] = 0
        if n == 0:
            return 0
        rows, valid, bounds, fl = self._str_inputs(off, data, validity, n, boff, bdata, bflags, nb)
        d = np.zeros(n, dtype=np.uint32)
        for i in range(n):
            cnt = 0
            for b in range(nb):
                none = bool(fl[b] & 4)
                if not desc:
                    right = True if not valid[i] else (False if none else rows[i] >= bounds[b])
                else:
                    right = True if (none or not valid[i]) else rows[i] < bounds[b]
                if right:
                    cnt += 1
                else:
                    break
            d[i] = cnt
        _arr(dest, np.uint32, n)[:] = d
        if _addr(counts):
            _arr(counts, np.uint64, nb + 1)[:] = np.bincount(d, minlength=nb + 1).astype(np.uint64)
        return 0
