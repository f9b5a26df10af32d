"""TEST INFRASTRUCTURE -- CPU restatement of the product's *interleaved* rANS container
(dcvc_b200/csrc/rans.cu, mode DCVC_RANS_INTERLEAVED).  Pure Python, small inputs only.

This format is the product's own (the reference has a single sequential coder per stream part,
DCVC-DC/src/cpp/py_rans/rans.cpp:76-168), so there is no reference output to pin it to: what is
pinned is (a) the symbol -> operations expansion, which is the reference's (rans.cpp:76-139:
interval of cdf[value]; out-of-support values escape through the last table entry followed by
4-bit bypass nibbles: a unary-in-15s digit count, then the digits of 2*|v| (+1 for negative)),
shared with the compat coder that IS byte-identical to the reference, and (b) losslessness.

Container: byte 0 = ((W-1) << 4) | 0xF (W <= 16) or byte 0 = 0x0E followed by W-1 as a
little-endian uint16 (16 < W <= 256); W-1 little-endian uint32 payload sizes; per warp payload:
32 little-endian uint32 lane states, then uint16 words in decode order.  A call of n symbols is
cut into W contiguous pieces (n // W each, the last takes the remainder); within a piece symbol k
belongs to lane k % 32, round k // 32.  Every lane is a 32-bit rANS state in [2^16, 2^32) with
16-bit renormalisation; in a round, operation slot s (0 = the interval, >= 1 = escape nibbles) is
coded for all lanes that have it, and the lanes that renormalise in that slot share consecutive
words of the warp's stream in lane order.
"""
import struct

import numpy as np

L32 = 1 << 16
PRECISION = 16


def part_range(n, parts, i):
    each = n // parts
    return i * each, (each if i < parts - 1 else n - each * (parts - 1))


def expand_symbol(symbol, index, cdf, sizes, offsets):
    """[(start, freq, bypass)] in decode order."""
    row = cdf[index]
    max_value = int(sizes[index]) - 2
    value = int(symbol) - int(offsets[index])
    raw = 0
    if value < 0:
        raw = -2 * value - 1
        value = max_value
    elif value >= max_value:
        raw = 2 * (value - max_value)
        value = max_value
    ops = [(int(row[value]), int(row[value + 1]) - int(row[value]), False)]
    if value == max_value:
        nb = 0
        while (raw >> (4 * nb)) != 0:
            nb += 1
        val = nb
        while val >= 15:
            ops.append((15, 0, True))
            val -= 15
        ops.append((val, 0, True))
        for j in range(nb):
            ops.append(((raw >> (4 * j)) & 15, 0, True))
    return ops


def encode(calls, tables, warps):
    """calls: [(symbols int16[n], indexes[n])] in push order -> container bytes."""
    cdf, sizes, offsets = tables
    payloads = []
    for w in range(warps):
        x = [L32] * 32
        events = []                                   # word groups in emission order
        for sym, idx in reversed(calls):
            off, cur = part_range(len(sym), warps, w)
            for r in reversed(range((cur + 31) // 32)):
                ops = []
                for lane in range(32):
                    k = r * 32 + lane
                    ops.append(expand_symbol(sym[off + k], idx[off + k], cdf, sizes, offsets)
                               if k < cur else [])
                for s in reversed(range(max(len(o) for o in ops))):
                    ev = []
                    for lane in range(32):
                        if s >= len(ops[lane]):
                            continue
                        start, freq, bypass = ops[lane][s]
                        x_max = (1 << 28) if bypass else (freq << 16)
                        if x[lane] >= x_max:
                            ev.append(x[lane] & 0xFFFF)
                            x[lane] >>= 16
                        if bypass:
                            x[lane] = (x[lane] << 4) | start
                        else:
                            x[lane] = ((x[lane] // freq) << PRECISION) + (x[lane] % freq) + start
                    if ev:
                        events.append(ev)
        words = []
        for ev in reversed(events):
            words.extend(ev)
        payloads.append(struct.pack("<32I", *x) + struct.pack(f"<{len(words)}H", *words))
    if warps > 16:      # the warp count does not fit the nibble: extended header
        head = bytes([0x0E]) + struct.pack("<H", warps - 1)
    else:
        head = bytes([((warps - 1) << 4) | 0xF])
    head += b"".join(struct.pack("<I", len(p)) for p in payloads[:-1])
    return head + b"".join(payloads)


def decode(stream, index_calls, tables):
    """index_calls: [indexes[n]] in pull order -> [int16[n]]."""
    cdf, sizes, offsets = tables
    if stream[0] == 0x0E:
        warps = struct.unpack_from("<H", stream, 1)[0] + 1
        pos = 3
    else:
        warps = (stream[0] >> 4) + 1
        assert stream[0] & 0xF == 0xF
        pos = 1
    lens = []
    for _ in range(warps - 1):
        lens.append(struct.unpack_from("<I", stream, pos)[0])
        pos += 4
    lens.append(len(stream) - pos - sum(lens))
    state = []
    for n in lens:
        p = stream[pos:pos + n]
        pos += n
        state.append({"x": list(struct.unpack_from("<32I", p, 0)),
                      "w": np.frombuffer(p, dtype="<u2", offset=128), "rp": 0})
    outs = []
    for idx in index_calls:
        out = np.zeros(len(idx), dtype=np.int16)
        for w, st in enumerate(state):
            off, cur = part_range(len(idx), warps, w)
            x = st["x"]
            for r in range((cur + 31) // 32):
                lanes = [lane for lane in range(32) if r * 32 + lane < cur]
                info = {}
                # slot 0: the interval
                for lane in lanes:
                    i = int(idx[off + r * 32 + lane])
                    row, size = cdf[i], int(sizes[i])
                    cum = x[lane] & 0xFFFF
                    s = int(np.searchsorted(row[:size], cum, side="right")) - 1
                    x[lane] = (int(row[s + 1]) - int(row[s])) * (x[lane] >> 16) + cum - int(row[s])
                    info[lane] = {"value": s, "max": size - 2, "off": int(offsets[i]),
                                  "phase": 1 if s == size - 2 else 0, "nb": 0, "raw": 0, "j": 0}
                for lane in lanes:
                    if x[lane] < L32:
                        x[lane] = (x[lane] << 16) | int(st["w"][st["rp"]])
                        st["rp"] += 1
                # escape slots
                while any(v["phase"] for v in info.values()):
                    vals = {}
                    for lane in lanes:
                        if info[lane]["phase"]:
                            vals[lane] = x[lane] & 15
                            x[lane] >>= 4
                    for lane in lanes:
                        if lane in vals and x[lane] < L32:
                            x[lane] = (x[lane] << 16) | int(st["w"][st["rp"]])
                            st["rp"] += 1
                    for lane, val in vals.items():
                        v = info[lane]
                        done = False
                        if v["phase"] == 1:
                            v["nb"] += val
                            if val != 15:
                                if v["nb"] == 0:
                                    done = True
                                else:
                                    v["phase"] = 2
                        else:
                            v["raw"] |= val << (4 * v["j"])
                            v["j"] += 1
                            done = v["j"] == v["nb"]
                        if done:
                            value = v["raw"] >> 1
                            v["value"] = -value - 1 if v["raw"] & 1 else value + v["max"]
                            v["phase"] = 0
                for lane in lanes:
                    out[off + r * 32 + lane] = info[lane]["value"] + info[lane]["off"]
        outs.append(out)
    return outs
