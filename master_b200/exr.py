"""The reference's checkpoint image format, Python mirror of master_b200/host/exr_io.hpp.

haste::save_exr / load_exr (exr.cpp:68-290) keep a render as a single-part scanline OpenEXR file: FLOAT channels R, G, B and
`denom` (the per-pixel sample count; the accumulators are stored UN-normalised so `master continue` / `master merge` can add
to them), file rows top-down while buffers are y-up, and every options.* / statistics.* / records[i].* entry as a string
attribute (Options.cpp:1185-1232, statistics.cpp:118-175).  No OpenEXR library is needed: the writer emits uncompressed
scanline blocks, the reader also takes ZIP / ZIPS blocks (OpenEXR's default when the reference writes).
"""
import struct
import zlib

import numpy as np

MAGIC = b"\x76\x2f\x31\x01"


def _attr(name, typ, value):
    return name.encode() + b"\0" + typ.encode() + b"\0" + struct.pack("<i", len(value)) + value


def save_exr(path, metadata, data):
    """data: (height, width, 3 or 4) float array, y-up, un-normalised sums (+ sample count) — exr.cpp:183-230."""
    data = np.asarray(data, dtype=np.float32)
    h, w, comps = data.shape
    assert comps in (3, 4) and h > 0 and w > 0
    names = ["B", "G", "R", "denom"][:comps]
    comp_of = [2, 1, 0, 3][:comps]
    longest = max([5] + [len(k) for k in metadata])
    out = [MAGIC, struct.pack("<i", 2 | (0x400 if longest > 31 else 0))]
    chlist = b"".join(n.encode() + b"\0" + struct.pack("<iB3xii", 2, 0, 1, 1) for n in names) + b"\0"
    window = struct.pack("<4i", 0, 0, w - 1, h - 1)
    out += [_attr("channels", "chlist", chlist), _attr("compression", "compression", b"\0"),
            _attr("dataWindow", "box2i", window), _attr("displayWindow", "box2i", window),
            _attr("lineOrder", "lineOrder", b"\0"), _attr("pixelAspectRatio", "float", struct.pack("<f", 1.0)),
            _attr("screenWindowCenter", "v2f", struct.pack("<2f", 0.0, 0.0)),
            _attr("screenWindowWidth", "float", struct.pack("<f", 1.0))]
    for k in sorted(metadata):
        out.append(_attr(k, "string", str(metadata[k]).encode()))
    out.append(b"\0")
    header = b"".join(out)
    row_bytes = comps * w * 4
    first = len(header) + h * 8
    offsets = np.arange(h, dtype=np.uint64) * np.uint64(8 + row_bytes) + np.uint64(first)
    flipped = data[::-1]  # file row y = buffer row height - y - 1 (exr.cpp:201-208)
    rows = np.ascontiguousarray(flipped[:, :, comp_of].transpose(0, 2, 1))  # (h, channel, w)
    with open(path, "wb") as f:
        f.write(header)
        f.write(offsets.astype("<u8").tobytes())
        for y in range(h):
            f.write(struct.pack("<ii", y, row_bytes))
            f.write(rows[y].astype("<f4").tobytes())


def _cstr(buf, pos):
    end = buf.index(b"\0", pos)
    return buf[pos:end].decode("latin-1"), end + 1


def load_exr(path):
    """-> (metadata dict, (height, width, 4) float32 array y-up: R, G, B, denom as stored; denom = 1 if the file has none,
    has_denom).  haste::load_exr for vec4 data (exr.cpp:244-290)."""
    buf = open(path, "rb").read()
    if buf[:4] != MAGIC:
        raise ValueError("%s: not an OpenEXR file" % path)
    version, = struct.unpack_from("<i", buf, 4)
    if version & 0x1A00:
        raise ValueError("tiled, deep and multi-part files are not what the renderer writes")
    pos, chans, compression, dw, metadata = 8, [], 0, None, {}
    while True:
        name, pos = _cstr(buf, pos)
        if not name:
            break
        typ, pos = _cstr(buf, pos)
        size, = struct.unpack_from("<i", buf, pos)
        pos += 4
        val = buf[pos:pos + size]
        pos += size
        if name == "channels":
            p = 0
            while val[p:p + 1] != b"\0":
                cn, p = _cstr(val, p)
                ptype, _, xs, ys = struct.unpack_from("<iB3xii", val, p)
                p += 16
                assert xs == 1 and ys == 1, "subsampled channels"
                chans.append((cn, ptype))
        elif name == "compression":
            compression = val[0]
        elif name == "dataWindow":
            dw = struct.unpack("<4i", val)
        elif typ == "string":
            metadata[name] = val.decode("latin-1")
    if compression not in (0, 2, 3):
        raise ValueError("only NONE / ZIPS / ZIP compression is read")
    w, h = dw[2] - dw[0] + 1, dw[3] - dw[1] + 1
    out = np.zeros((h, w, 4), np.float32)
    out[:, :, 3] = 1.0
    has_denom = False
    line_bytes = sum(w * (2 if t == 1 else 4) for _, t in chans)
    per_block = 16 if compression == 3 else 1
    blocks = (h + per_block - 1) // per_block
    offs = struct.unpack_from("<%dQ" % blocks, buf, pos)
    for o in offs:
        y0, csize = struct.unpack_from("<ii", buf, o)
        y0 -= dw[1]
        lines = min(per_block, h - y0)
        want = lines * line_bytes
        chunk = buf[o + 8:o + 8 + csize]
        if compression != 0 and csize != want:
            t = np.frombuffer(zlib.decompress(chunk), dtype=np.uint8).astype(np.int64)
            t[1:] -= 128
            t = (np.cumsum(t) & 0xFF).astype(np.uint8)  # undo the byte predictor
            half = (want + 1) // 2
            raw = np.empty(want, np.uint8)
            raw[0::2], raw[1::2] = t[:half], t[half:]  # undo the even / odd byte split
            chunk = raw.tobytes()
        p = 0
        for l in range(lines):
            row = h - 1 - (y0 + l)
            for cn, t in chans:
                nb = w * (2 if t == 1 else 4)
                comp = {"R": 0, "G": 1, "B": 2, "denom": 3}.get(cn)
                if comp is not None:
                    assert t == 2, "channel %s is not FLOAT" % cn
                    out[row, :, comp] = np.frombuffer(chunk, "<f4", w, p)
                    has_denom |= comp == 3
                p += nb
    return metadata, out, has_denom


# ---------------------------------------------------------------- metadata (Options::to_dict, statistics_t::to_dict)
def _f(x):
    return "%f" % float(x)  # std::to_string(float / double)


def options_to_dict(technique="BPT", input0="", output="", width=512, height=512, num_samples=0, num_seconds=0.0, radius=0.1,
                    alpha=0.75, beta=1.0, roulette=0.9, lights=1.0, num_photons=0, max_path=2 ** 63 - 1, enable_vc=True,
                    enable_vm=True, from_light=True, seed=0, enable_seed=False, camera_id=0, num_threads=0, snapshot=0,
                    sky_horizon=(0, 0, 0), sky_zenith=(0, 0, 0)):
    """Options::to_dict (Options.cpp:1185-1232), same keys and std::to_string formatting."""
    b = lambda v: str(int(bool(v)))
    return {
        "options.input0": input0, "options.input1": "", "options.output": output, "options.reference": "",
        "options.technique": technique, "options.action": "Render", "options.num_photons": str(int(num_photons)),
        "options.radius": _f(radius), "options.max_path": str(int(max_path)), "options.alpha": _f(alpha), "options.beta": _f(beta),
        "options.roulette": _f(roulette), "options.batch": "1", "options.quiet": "0", "options.enable_vc": b(enable_vc),
        "options.enable_vm": b(enable_vm), "options.from_light": b(from_light), "options.lights": _f(lights),
        "options.num_samples": str(int(num_samples)), "options.num_seconds": _f(num_seconds),
        "options.num_threads": str(int(num_threads)), "options.reload": "0", "options.enable_seed": b(enable_seed),
        "options.enable_ui": "0", "options.seed": str(int(seed)), "options.snapshot": str(int(snapshot)),
        "options.camera_id": str(int(camera_id)), "options.width": str(int(width)), "options.height": str(int(height)),
        "options.sky_horizon": "x".join(_f(v) for v in sky_horizon), "options.sky_zenith": "x".join(_f(v) for v in sky_zenith),
    }


STAT_COUNTERS = ("num_samples", "num_basic_rays", "num_shadow_rays", "num_tentative_rays", "num_photons", "num_scattered")
STAT_TIMES = ("total_time", "scatter_time", "build_time", "gather_time", "merge_time", "density_time", "intersect_time",
              "trace_eye_time", "trace_light_time")


def statistics_to_dict(records=(), **values):
    """statistics_t::to_dict (statistics.cpp:118-175); records: iterable of dicts with sample_index, rms_error, abs_error,
    clock_time, frame_duration, numeric_errors."""
    d = {}
    for k in STAT_COUNTERS:
        d["statistics." + k] = str(int(values.get(k, 0)))
    for k in STAT_TIMES:
        d["statistics." + k] = _f(values.get(k, 0.0))
    for r in records:
        i = int(r["sample_index"])
        for k in ("rms_error", "abs_error", "clock_time", "frame_duration"):
            d["records[%d].%s" % (i, k)] = _f(r.get(k, 0.0))
        d["records[%d].numeric_errors" % i] = str(int(r.get("numeric_errors", 0)))
    return d


def merge_exr(dst, fst, snd):
    """`master merge` (merge_exr, Options.cpp:1340-1409): pixel sums and sample counts add, counters and timers add, the
    records collapse to one whose frame_duration is the total rendering time.  This is what a multi-GPU render is: one file
    per rank's samples, merged."""
    m1, a, _ = load_exr(fst)
    m2, b, _ = load_exr(snd)
    if a.shape != b.shape:
        raise ValueError("Sizes of '%s' and '%s' doesn't match." % (fst, snd))
    if m1.get("options.technique") != m2.get("options.technique"):
        raise ValueError("Cannot merge images rendered using different techniques.")
    out = dict(m1)
    out["options.output"] = dst
    for k in STAT_COUNTERS:
        out["statistics." + k] = str(int(m1.get("statistics." + k, 0)) + int(m2.get("statistics." + k, 0)))
    for k in STAT_TIMES:
        out["statistics." + k] = _f(float(m1.get("statistics." + k, 0)) + float(m2.get("statistics." + k, 0)))
    duration = sum(float(v) for m in (m1, m2) for k, v in m.items() if k.startswith("records[") and k.endswith(".frame_duration"))
    rec = sorted({int(k[8:k.index("]")]) for k in m1 if k.startswith("records[")})
    out = {k: v for k, v in out.items() if not k.startswith(("records[", "measurements["))}
    if rec:
        last = rec[-1]
        for k, v in m1.items():
            if k.startswith("records[%d]." % last):
                out[k] = v
        out["records[%d].frame_duration" % last] = _f(duration)
    save_exr(dst, out, a + b)
    return out
