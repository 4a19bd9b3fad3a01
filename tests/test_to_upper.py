"""strings.ToUpper as Go 1.19 computes it -- what checkGPUtype / checkDCUtype / checkMLUtype fold both sides of their
substring test with (/root/reference/pkg/device/nvidia/device.go:66-86, hygon/device.go:49-69,
cambricon/device.go:63-83; go.mod: go 1.19 => Unicode 13.0.0 tables).  The host mirror (vgs_host_to_upper, used by
vgs_host_typemask*) and the oracle (ref_strings_to_upper, used by ref_checkType) carry their own decoders and their own
copy of the generated table; both are held against known answers, against this interpreter's Unicode data for every code
point, and against a restatement of strings.Map written here.  No GPU."""
import unicodedata

import numpy as np
import pytest

import oracle
import vgpu_b200 as v

IMPLS = [("host mirror", v.capi.host_to_upper), ("oracle", oracle.to_upper)]
FFFD = "�".encode()

# case pairs that entered Unicode after 13.0.0 (all in 14.0.0): not in Go 1.19's tables
LATER = {0x2C5F, 0xA7C1, 0xA7D1, 0xA7D7, 0xA7D9} | set(range(0x10597, 0x105BD))

# written down from the Unicode character database and the Go sources, not computed:
KNOWN = [
    (b"", b""),
    (b"nvidia-tesla v100", b"NVIDIA-TESLA V100"),
    ("a100,ß".encode(), "A100,ß".encode()),              # U+00DF has no simple uppercase (the full one, SS, is not used)
    ("ǆ ǅ Ǆ".encode(), "Ǆ Ǆ Ǆ".encode()),                # U+01C6 / U+01C5 (title case) -> U+01C4
    ("nvıdıa".encode(), b"NVIDIA"),                        # U+0131 dotless i -> ASCII I
    ("teſla".encode(), b"TESLA"),                          # U+017F long s -> ASCII S
    ("µ".encode(), "Μ".encode()),                          # U+00B5 micro sign -> U+039C
    ("ÿ".encode(), "Ÿ".encode()),                          # U+00FF -> U+0178
    ("ᾳ".encode(), "ᾼ".encode()),                          # U+1FB3 -> U+1FBC (full mapping would be two letters)
    ("ŉ ǰ ΐ ﬁ ﬓ".encode(), "ŉ ǰ ΐ ﬁ ﬓ".encode()),          # full mappings only: unchanged
    ("ა".encode(), "Ა".encode()),                          # Georgian Mkhedruli -> Mtavruli (Unicode 11)
    ("ⱟ ꟁ 𐖗".encode(), "ⱟ ꟁ 𐖗".encode()),                # pairs of Unicode 14.0.0: unknown to Go 1.19
    ("ɐ".encode(), "Ɐ".encode()),                          # 2-byte encoding grows to 3
    ("𐐨".encode(), "𐐀".encode()),                        # Deseret, 4-byte encodings
    ("İ K".encode(), "İ K".encode()),                      # already upper case (U+0130, U+212A Kelvin sign)
    # bytes outside a valid encoding: strings.Map ranges over the string, each such byte is U+FFFD
    (b"\xff", FFFD),
    (b"a\xc3", b"A" + FFFD),                               # truncated
    (b"\xe2\x82", FFFD * 2),                               # truncated 3-byte form: one U+FFFD per byte
    (b"\xed\xa0\x80", FFFD * 3),                           # a surrogate is not a valid encoding
    (b"\xc0\xaf", FFFD * 2),                               # overlong
    (b"\xf4\x90\x80\x80x", FFFD * 4 + b"X"),               # above U+10FFFF
    (FFFD + b"q", FFFD + b"Q"),                            # a real U+FFFD stays what it is
    (b"tesla\x80", b"TESLA" + FFFD),
]


def simple_upper(cp: int) -> int:
    """Simple_Uppercase_Mapping of Unicode 13.0.0 from this interpreter's data (see tools/make_upper_table.py)"""
    if cp in LATER:
        return cp
    c = chr(cp)
    u = c.upper()
    if len(u) == 1:
        return ord(u)
    t = c.title()
    return ord(t) if len(t) == 1 and t != c else cp


def go_to_upper(s: bytes) -> bytes:
    """strings.ToUpper restated: ASCII fast path, else rune by rune with (U+FFFD, 1) for undecodable bytes"""
    if all(b < 0x80 for b in s):
        return s.upper()
    out, i = bytearray(), 0
    while i < len(s):
        b0 = s[i]
        need = 1 if b0 < 0x80 else 2 if 0xC2 <= b0 <= 0xDF else 3 if 0xE0 <= b0 <= 0xEF else 4 if 0xF0 <= b0 <= 0xF4 else 0
        try:
            ch = s[i:i + need].decode("utf-8") if need else None  # strict: no overlong forms, no surrogates
        except UnicodeDecodeError:
            ch = None
        if ch is None or len(ch) != 1:
            out += FFFD
            i += 1
        else:
            out += chr(simple_upper(ord(ch))).encode()
            i += need
    return bytes(out)


@pytest.mark.parametrize("name,fn", IMPLS, ids=[n for n, _ in IMPLS])
def test_known_answers(built, name, fn):
    for s, want in KNOWN:
        assert fn(s) == want, (name, s)
        assert go_to_upper(s) == want, ("restatement", s)


@pytest.mark.parametrize("name,fn", IMPLS, ids=[n for n, _ in IMPLS])
def test_every_code_point(built, name, fn):
    """one call over all 1 112 064 scalar values, checked against this interpreter's Unicode data (15.0.0) cut back
    to 13.0.0; the generated table holds 1384 non-ASCII pairs"""
    assert unicodedata.unidata_version == "15.0.0"
    cps = [cp for cp in range(1, 0x110000) if not 0xD800 <= cp <= 0xDFFF]
    src = "".join(map(chr, cps)).encode()
    want = "".join(chr(simple_upper(cp)) for cp in cps)
    got = fn(src).decode("utf-8")
    assert len(got) == len(want)
    if got != want:
        bad = [(hex(cp), hex(ord(g)), hex(ord(w))) for cp, g, w in zip(cps, got, want) if g != w]
        raise AssertionError((name, len(bad), bad[:10]))
    assert sum(1 for cp, w in zip(cps, want) if cp >= 0x80 and ord(w) != cp) == 1384


def test_random_bytes_agree(built):
    """random byte strings (mostly invalid UTF-8) and random scalar values: mirror == oracle == restatement"""
    rng = np.random.default_rng(31)
    for _ in range(3000):
        n = int(rng.integers(0, 24))
        kind = rng.random()
        if kind < 0.4:
            s = bytes(int(x) for x in rng.integers(1, 256, n))
        elif kind < 0.7:
            s = "".join(chr(int(c)) for c in rng.choice([0x61, 0xDF, 0x131, 0x17F, 0x3B1, 0x1F80, 0x2C5F, 0x10428, 0x1E943, 0x7A], n)).encode()
            cut = int(rng.integers(0, len(s) + 1))
            s = s[:cut] + bytes([int(rng.integers(0x80, 0x100))]) + s[cut:]
        else:
            pts = [int(c) for c in rng.integers(1, 0x110000, n)]
            s = "".join(chr(c) for c in pts if not 0xD800 <= c <= 0xDFFF).encode()
        want = go_to_upper(s)
        assert v.capi.host_to_upper(s) == want, s
        assert oracle.to_upper(s) == want, s


def test_type_policy_folds_like_go(built):
    """through checkType: an annotation written with U+0131 / U+017F names an ASCII card, U+00DF does not; Greek and
    Cyrillic card names fold; the mirror's mask equals the oracle's on every combination"""
    cards = ["NVIDIA-Tesla V100", "NVIDIA-NVIDIA A100", "NVIDIA-Ωmega αβγ", "NVIDIA-карта Ж1", "NVIDIA-STRASSE", "NVIDIA-straße"]
    cases = [
        ("nvıdıa", 0b111111), ("teſla", 0b000001), ("ΩMEGA ΑΒΓ", 0b000100), ("ωmega", 0b000100), ("КАРТА ж1", 0b001000),
        ("strasse", 0b010000), ("STRAßE", 0b100000), ("straẞe", 0b000000),  # U+1E9E is upper case already, not U+00DF
        ("a100,карта", 0b001010), ("ᾀ", 0), ("", 0b111111),
    ]
    for tok, want in cases:
        an = {v.capi.ANNO_USE: tok}
        m = v.capi.host_typemask("NVIDIA", an, cards)
        assert m == want, (tok, bin(m))
        assert oracle.typemask(an, "NVIDIA", cards)[0] == want, tok
        an = {v.capi.ANNO_NOUSE: tok}
        assert v.capi.host_typemask("NVIDIA", an, cards) == (0b111111 & ~want), tok
        assert oracle.typemask(an, "NVIDIA", cards)[0] == (0b111111 & ~want), tok
    mlu = ["MLU370-Ж", "MLU290-ж"]
    for tok in ("ж", "Ж", "mlu370-ж"):
        an = {v.capi.ANNO_MLU_USE: tok}
        assert v.capi.host_typemask_mlu(0, an, mlu)[0] == oracle.typemask(an, "MLU", mlu)[0] == (0b11 if len(tok) <= 2 else 0b01), tok


def test_committed_tables_are_what_the_generator_writes(tmp_path):
    """csrc/vgs_upper_table.inc and oracle/upper_table.inc are generated files: byte for byte what
    tools/make_upper_table.py writes from this interpreter's Unicode data, and equal to each other"""
    import importlib.util
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("make_upper_table", os.path.join(root, "tools", "make_upper_table.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    pairs = gen.table()
    out = tmp_path / "t.inc"
    gen.emit(str(out), gen.runs(pairs), len(pairs))
    want = out.read_bytes()
    for rel in ("4paradigm-k8s-vgpu-scheduler_b200/csrc/vgs_upper_table.inc", "oracle/upper_table.inc"):
        assert open(os.path.join(root, rel), "rb").read() == want, rel
