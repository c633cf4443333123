"""The device's JSON line scanner, compiled for the CPU (tests/json_scan_harness.cpp), against Python's json module:
field spans on the fixture lines, on hand-written edge cases and on randomly nested documents, at every alignment of
the line inside the text buffer (the scanner works on aligned 32-byte blocks)."""
import ctypes
import json
import os
import random
import subprocess

import numpy as np
import pytest

import fake_kubernetes as fk
from oracle import c_oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def scan(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("jsonscan") / "libjsonscan.so")
    subprocess.run(["g++", "-O1", "-shared", "-fPIC", "-o", so, os.path.join(ROOT, "tests", "json_scan_harness.cpp")], check=True)
    lib = ctypes.CDLL(so)
    lib.scan_line_host.argtypes = [ctypes.c_char_p, ctypes.c_ulonglong, ctypes.c_ulonglong, ctypes.POINTER(ctypes.c_longlong)]
    lib.scan_line_host.restype = ctypes.c_int

    def run(line: bytes, shift: int = 0):
        buf = b"#" * shift + line + b"\xff" * 64           # garbage behind the line must not be looked at
        out = (ctypes.c_longlong * 11)()
        ok = lib.scan_line_host(buf, shift, shift + len(line), out)
        vals = []
        for k in range(5):
            s, n = out[2 * k], out[2 * k + 1]
            vals.append(None if s < 0 else buf[s:s + n])
        return bool(ok), vals, bool(out[10])
    return run


def expected(line: bytes):
    """What the fields are according to json.loads (raw value bytes, so only for values without escapes)."""
    try:
        ev = json.loads(line)
    except ValueError:
        return None
    if not isinstance(ev, (dict, list)):
        return None

    def s(v):
        return v.encode() if isinstance(v, str) else None
    typ = uid = ns = rv = ph = None
    has_status = False
    if isinstance(ev, dict):
        typ = s(ev.get("type"))
        obj = ev.get("object")
        if isinstance(obj, dict):
            md = obj.get("metadata")
            if isinstance(md, dict):
                uid, ns, rv = s(md.get("uid")), s(md.get("namespace")), s(md.get("resourceVersion"))
            st = obj.get("status")
            if isinstance(st, dict):
                has_status = True
                ph = s(st.get("phase"))
    return [typ, uid, ns, rv, ph], has_status


def check(scan, line: bytes, shifts=(0, 1, 7, 31, 32, 45)):
    want = expected(line)
    for sh in shifts:
        ok, vals, hs = scan(line, sh)
        if want is None:
            continue            # what the scanner does with malformed lines is pinned by the named cases below
        assert ok, (line, sh)
        assert (vals, hs) == (want[0], want[1]), (line, sh)


def test_fixture_lines(scan):
    cols = c_oracle.generate(7, 3000, 200, dist=1, p_del=0.08, p_nostatus=0.05, p_null=0.05)
    text = fk.watch_json_lines(fk.events_from_columns(cols), seed=3)
    lines = [l for l in text.split(b"\n") if l]
    assert len(lines) == 3000
    for i, line in enumerate(lines):
        check(scan, line, shifts=(i % 64,))


EDGE = [
    b'{"type":"ADDED","object":{"metadata":{"uid":"u","namespace":"n","resourceVersion":"5"},"status":{"phase":"Running"}}}',
    b'  {  "type" : "ADDED" , "object" : { "metadata" : { "uid" : "u" } , "status" : null } }  ',
    b'{"object":{"status":{},"metadata":{"namespace":null,"uid":5,"resourceVersion":["7"]}},"type":"DELETED"}',
    b'{"type":"MODIFIED","object":{"metadata":{"uid":"a","uid":"b"},"status":{"phase":"x","phase":null}}}',
    b'{"type":"ADDED","other":{"metadata":{"uid":"decoy"}},"object":{"spec":{"metadata":{"uid":"decoy2"}},"metadata":{"uid":"real"}}}',
    b'{"type":"ADDED","object":{"metadata":[{"uid":"decoy"}],"status":[{"phase":"decoy"}]}}',
    b'{"type":"ADDED","object":[{"metadata":{"uid":"decoy"}}]}',
    b'[{"type":"ADDED"}]',
    b'{"type":"ADDED","object":{"metadata":{"annotations":{"a":"x\\"y\\\\","b":"{[:\\\\\\"]}"},"uid":"u2","labels":{"uid":"no"}}}}',
    b'{"a":"\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\\","type":"ADDED","object":{"metadata":{"uid":"after-a-long-backslash-run"}}}',
    b'{"type":"ADDED","object":{"metadata":{"uid":"' + b"x" * 100 + b'","namespace":"' + b"y" * 70 + b'"}}}',
    b'{"type":"ADDED","object":{"metadata":{"uid":                                             "far"}}}',
    b'{"type":"ADDED","object":{"metadata":{"labels":{"a":{"b":{"c":{"d":{"e":[[[[{"uid":"deep"}]]]]}}}}},"uid":"shallow"}}}',
    b'{"type":"ADDED","object":{"metadata":{"uid":"u"},"status":{"phase":"Running","conditions":[{"type":"Ready","phase":"no"}]}}}',
    # more look-alike keys than the scanner's candidate queue holds (8): keys of the lengths it looks for, before, between and after the real ones
    ('{"type":"ADDED","object":{"metadata":{' + ",".join('"k%02d":"v%d"' % (k, k) for k in range(30)) + ',"uid":"u-mid",'
     + ",".join('"z%02d":%d' % (k, k) for k in range(30)) + ',"namespace":"ns-late","uid":"u-last"},"status":{'
     + ",".join('"p%04d":"x"' % k for k in range(20)) + ',"phase":"Running"}}}').encode(),
    b'{"type":"ADDED","object":{"metadata":{"uid":5,"namespace":"x","resourceVersion":null},"status":{"phase":7,"other":"Running"}}}',
    b'{"type":"ADDED","object":{"metadata":{"uid":"first","uid":7,"namespace":"n1","namespace":"n2"}}}',
    b'{"type":"ADDED","object":{"metadata":{"uid":null}},"type":"DELETED"}',
    b'{"type" :   "MODIFIED"  ,"object":{"metadata":{"uid"\t:\t"tabbed","namespace":' + b" " * 70 + b'"after-seventy-spaces"}}}',
    b'{}', b'[]', b'{"type":""}', b'{"type":"ADDED","object":{"metadata":{"uid":""}}}',
    '{"type":"ADDED","object":{"metadata":{"uid":"u","namespace":"é中\U0001f600"}}}'.encode(),
]


@pytest.mark.parametrize("k", range(len(EDGE)))
def test_edge_cases(scan, k):
    assert expected(EDGE[k]) is not None
    check(scan, EDGE[k], shifts=range(0, 70))


@pytest.mark.parametrize("line", [
    b'', b'   ', b'{"type":"ADDED"', b'{"type":"ADDED"}}', b'}{', b'"type"', b'x{}', b'{"a":"unterminated}',
    b'{"type":"ADDED","object":{"metadata":{"uid":"u"}', b'{"a":"b\\"}',
])
def test_lines_that_are_not_one_container_are_refused(scan, line):
    for sh in (0, 5, 31):
        assert scan(line, sh)[0] is False


def _random_doc(rnd, depth=0):
    names = ["type", "object", "metadata", "status", "uid", "namespace", "resourceVersion", "phase", "spec", "x", "labels"]
    nasty = ['a"b', "back\\slash", "{[:,]}", "\\", '\\"', "é", "\n\t", "", "plain", ':"', "\\\\\\"]

    def value(d):
        r = rnd.random()
        if d < 5 and r < 0.35:
            return _random_doc(rnd, d + 1)
        if d < 5 and r < 0.45:
            return [value(d + 1) for _ in range(rnd.randrange(0, 3))]
        if r < 0.55:
            return rnd.choice([None, True, False, 5, -1.5e3])
        if r < 0.8:
            return rnd.choice(["ADDED", "u-%d" % rnd.randrange(100), "ns", "12", "Running", ""])     # safe: no escapes needed
        return rnd.choice(nasty)
    doc = {}
    for _ in range(rnd.randrange(0, 6)):
        k = rnd.choice(names) if rnd.random() < 0.85 else rnd.choice(nasty)
        v = value(depth)
        if k in ("type", "uid", "namespace", "resourceVersion", "phase") and isinstance(v, str) and v in nasty:
            v = "safe"
        doc[k] = v
    return doc


def test_random_documents(scan):
    rnd = random.Random(20251107)
    for i in range(4000):
        doc = _random_doc(rnd)
        if rnd.random() < 0.5:       # steer towards the interesting shape
            doc = {"type": "MODIFIED", "object": {"metadata": _random_doc(rnd, 2), "status": rnd.choice([None, _random_doc(rnd, 2)]),
                                                  "spec": _random_doc(rnd, 2)}, **doc}
        sep = rnd.choice([(",", ":"), (", ", ": "), (" ,  ", "  :   ")])
        line = json.dumps(doc, separators=sep, ensure_ascii=rnd.random() < 0.5).encode()
        check(scan, line, shifts=(i % 32, (i * 7) % 64))


def test_unescape_matches_json_loads(scan, tmp_path_factory):
    so = str(tmp_path_factory.mktemp("jsonscan2") / "libjsonscan.so")
    subprocess.run(["g++", "-O1", "-shared", "-fPIC", "-o", so, os.path.join(ROOT, "tests", "json_scan_harness.cpp")], check=True)
    lib = ctypes.CDLL(so)
    lib.unescape_host.argtypes = [ctypes.c_char_p, ctypes.c_ulonglong, ctypes.c_ulonglong, ctypes.c_char_p, ctypes.c_ulonglong]
    lib.unescape_host.restype = ctypes.c_longlong

    def dec(body: bytes):
        buf = b"xx" + body + b'"' + b"\xff" * 16
        out = ctypes.create_string_buffer(4 * len(body) + 8)
        n = lib.unescape_host(buf, 2, len(body), out, len(out))
        return None if n < 0 else out.raw[:n]
    rnd = random.Random(5)
    alphabet = ['a', 'Z', '"', '\\', '/', '\b', '\f', '\n', '\r', '\t', 'é', '中', '\U0001f600', '\x00', '\x7f', ' ', '߿', 'ࠀ', '￿']
    for i in range(3000):
        text = "".join(rnd.choice(alphabet) for _ in range(rnd.randrange(0, 12)))
        body = json.dumps(text, ensure_ascii=rnd.random() < 0.7).encode()[1:-1]
        if rnd.random() < 0.3:
            body = body.replace(b"/", b"\\/")
        assert dec(body) == text.encode(), body
    assert dec(b"\\u0041\\u00e9\\ud83d\\ude00") == "Aé\U0001f600".encode()
    for broken in (b"\\q", b"\\u12", b"\\u12zz", b"\\ud83d", b"\\ud83dx", b"\\ud83d\\u0041", b"\\ude00"):
        assert dec(broken) is None, broken
