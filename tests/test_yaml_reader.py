"""The labeler's YAML reader (labeler/yaml_lite.cpp, through the test helper tools/yaml_dump) against PyYAML on generated and
hand-written documents: anchors / aliases / merge keys, literal and folded block scalars with chomping indicators, tags,
directives, multi-document files (first document only, as yaml.v2's Unmarshal reads them).  The reference parses its config
file with sigs.k8s.io/yaml = gopkg.in/yaml.v2 (vendor/sigs.k8s.io/yaml/yaml.go:35-49,121); PyYAML is the independent YAML 1.1
implementation available here.  Where the two real implementations differ (a merge key placed AFTER an explicit key:
yaml.v2 decode.go:660-664 lets the merge overwrite) the expectation is written out by hand from the Go code."""
import json
import math
import os
import re
import subprocess

import pytest
import yaml
from hypothesis import HealthCheck, given, settings, strategies as st

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DUMP = os.environ.get("GFD_YAML_DUMP", os.path.join(ROOT, "tools", "yaml_dump"))  # GFD_YAML_DUMP=<ASan build> for sanitizer runs


def product_parse(text):
    p = subprocess.run([DUMP], input=text.encode(), capture_output=True, timeout=30)
    if p.returncode != 0:
        return ("error", p.stderr.decode().strip())

    def hook(d):
        if set(d) == {"__float__"}:
            return float(d["__float__"])
        return d
    v = json.loads(p.stdout.decode(), object_hook=hook)

    def fix(x):
        if isinstance(x, str) and x in ("__nan__", "__inf__", "__-inf__"):
            return {"__nan__": math.nan, "__inf__": math.inf, "__-inf__": -math.inf}[x]
        if isinstance(x, list):
            return [fix(i) for i in x]
        if isinstance(x, dict):
            return {k: fix(i) for k, i in x.items()}
        return x
    return ("ok", fix(v))


def same(a, b):
    if isinstance(a, float) and isinstance(b, float):
        return (math.isnan(a) and math.isnan(b)) or a == b
    if type(a) is not type(b):
        return False
    if isinstance(a, list):
        return len(a) == len(b) and all(same(x, y) for x, y in zip(a, b))
    if isinstance(a, dict):
        return list(a) == list(b) and all(same(a[k], b[k]) for k in a)
    return a == b


WORDS = st.text(alphabet="abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789-_./ ", min_size=1, max_size=14).filter(
    lambda s: s.strip() == s and s not in ("y", "Y", "n", "N"))  # yaml.v2 reads a bare y/Y/n/N as a bool (YAML 1.1), PyYAML as a string
LINES = st.lists(st.text(alphabet="abcdefghijklmnopqrstuvwxyz0123456789 -_.:#,[]{}&*!|>'\"%@", min_size=0, max_size=18), min_size=1, max_size=5)
MULTILINE = st.builds(lambda ls, tail: "\n".join(ls) + tail, LINES, st.sampled_from(["", "\n", "\n\n"]))
SCALAR = st.one_of(st.none(), st.booleans(), st.integers(-10 ** 6, 10 ** 12), st.sampled_from([0.5, -2.25, 1e9, 60.0]), WORDS, MULTILINE)
KEYS = st.text(alphabet="abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789_.-", min_size=1, max_size=10).filter(
    lambda k: not k[0] in "-." and k.lower() not in ("y", "n", "yes", "no", "on", "off", "true", "false", "null") and not k.replace(".", "").replace("_", "").isdigit()
    and not k[0].isdigit())
TREE = st.recursive(SCALAR, lambda kids: st.one_of(st.lists(kids, max_size=4), st.dictionaries(KEYS, kids, max_size=4)), max_leaves=12)


class _Literal(yaml.SafeDumper):
    pass


def _str_rep(style):
    def rep(dumper, data):
        return dumper.represent_scalar("tag:yaml.org,2002:str", data, style=style if "\n" in data else None)
    return rep


class _Folded(yaml.SafeDumper):
    pass


_Literal.add_representer(str, _str_rep("|"))
_Folded.add_representer(str, _str_rep(">"))


@st.composite
def shared_doc(draw):
    """A document in which one sub-tree object is referenced several times, so PyYAML emits &id001 / *id001."""
    shared = draw(st.one_of(st.dictionaries(KEYS, SCALAR, min_size=1, max_size=3), st.lists(SCALAR, min_size=1, max_size=3)))
    other = draw(TREE)
    shape = draw(st.sampled_from(["list", "map", "nested"]))
    if shape == "list":
        return [shared, other, shared, {"again": shared}]
    if shape == "map":
        return {"first": shared, "x": other, "second": shared}
    return {"outer": {"inner": [shared, shared]}, "tail": other, "last": shared}


@settings(max_examples=400, deadline=None, suppress_health_check=[HealthCheck.too_slow, HealthCheck.filter_too_much])
@given(doc=st.one_of(TREE, shared_doc()), style=st.sampled_from(["block", "block4", "literal", "folded", "flow_leaves", "explicit_start", "wrapped", "flow_wrapped"]))
def test_reader_agrees_with_pyyaml(doc, style):
    def has_yn(x):
        # Known differences between the two real implementations, not tested here (the hand-written cases below pin what yaml.v2
        # does): a bare y/Y/n/N is a bool for yaml.v2 and a string for PyYAML; number-like words PyYAML leaves as strings
        # (08, 0o17, 1e3, 0E0) are numbers under yaml.v2's ParseInt(base 0) / yamlStyleFloat rules.
        if isinstance(x, str):
            return x in ("y", "Y", "n", "N") or re.match(r"^[-+.]?[0-9]", x) is not None
        if isinstance(x, list):
            return any(has_yn(i) for i in x)
        if isinstance(x, dict):
            return any(has_yn(i) for i in x.values())
        return False
    if has_yn(doc):
        return
    kw = dict(width=100000, default_flow_style=False, sort_keys=False)
    dumper = yaml.SafeDumper
    if style == "block4":
        kw["indent"] = 4
    elif style == "literal":
        dumper = _Literal
    elif style == "folded":
        dumper = _Folded
    elif style == "flow_leaves":
        kw["default_flow_style"] = None
    elif style == "explicit_start":
        kw["explicit_start"] = True
        kw["version"] = (1, 1)
    elif style == "wrapped":          # narrow page: PyYAML continues quoted (and plain) scalars on following lines
        kw["width"] = 24
    elif style == "flow_wrapped":     # flow collections broken over several lines
        kw["width"] = 24
        kw["default_flow_style"] = True
    text = yaml.dump(doc, Dumper=dumper, **kw)
    try:
        want = yaml.safe_load(text)
    except yaml.YAMLError:
        return
    status, got = product_parse(text)
    assert status == "ok", (text, got)   # everything PyYAML's dumper emits for these documents is read
    assert same(got, want), (text, got, want)


HAND = [
    # merge keys: the usual form, a list of maps (earlier wins), and yaml.v2's "where it stands" rule
    ("a: &a {x: 1, w: 2}\nb:\n  <<: *a\n  w: 3\n", {"a": {"x": 1, "w": 2}, "b": {"x": 1, "w": 3}}),
    ("a: &a {x: 1}\nb: &b {x: 2, z: 9}\nc:\n  <<: [*a, *b]\n", {"a": {"x": 1}, "b": {"x": 2, "z": 9}, "c": {"x": 1, "z": 9}}),
    ("a: &a {w: 2}\nb:\n  w: 3\n  <<: *a\n", {"a": {"w": 2}, "b": {"w": 2}}),          # decode.go:660-664: the merge overwrites the earlier key
    ("a: &a {w: 2}\nb: {w: 3, <<: *a, z: 1}\n", {"a": {"w": 2}, "b": {"w": 2, "z": 1}}),
    ("a: &a {w: 2}\nb:\n  \"<<\": *a\n", {"a": {"w": 2}, "b": {"<<": {"w": 2}}}),      # a quoted "<<" is an ordinary key
    # anchors on scalars, sequences and nested blocks
    ("x: &s hello\nw: *s\nz: [*s, *s]\n", {"x": "hello", "w": "hello", "z": ["hello", "hello"]}),
    ("l: &l\n- 1\n- 2\nm: *l\n", {"l": [1, 2], "m": [1, 2]}),
    ("- &m\n  k: v\n- *m\n", [{"k": "v"}, {"k": "v"}]),
    # block scalars: clip / strip / keep, indentation indicator, blank and more-indented lines, '#' is content
    ("s: |\n  a\n  b\n", {"s": "a\nb\n"}),
    ("s: |-\n  a\n  b\n\n", {"s": "a\nb"}),
    ("s: |+\n  a\n\n\nt: 1\n", {"s": "a\n\n\n", "t": 1}),
    ("s: |2\n    two more\n  base\n", {"s": "  two more\nbase\n"}),
    ("s: |\n  # not a comment\n  x: y\nt: 2\n", {"s": "# not a comment\nx: y\n", "t": 2}),
    ("s: >\n  one\n  two\n\n  three\n    code\n  four\n", {"s": "one two\nthree\n  code\nfour\n"}),
    ("s: >-\n  one\n  two\n", {"s": "one two"}),
    ("- |\n  item\n- >-\n  folded\n  item\n", ["item\n", "folded item"]),
    ("s: |\n\nt: 1\n", {"s": "", "t": 1}),
    # quoted scalars and flow collections over several lines (line folding inside quotes; an escaped break joins)
    ("a: [1,\n  2, # c\n  {k: \"x\n     y\"}\n  ]\nb: \"one\n  two\n\n  three\\\n  four\"\nc: 's1\n  s2'\nd: ok\n",
     {"a": [1, 2, {"k": "x y"}], "b": "one two\nthreefour", "c": "s1 s2", "d": "ok"}),
    ("sharing:\n  timeSlicing:\n    resources: [\n      {name: nvidia.com/gpu,\n       replicas: 2}\n    ]\n",
     {"sharing": {"timeSlicing": {"resources": [{"name": "nvidia.com/gpu", "replicas": 2}]}}}),
    ("- \"a # not a comment\n  b\"\n- x\n", ["a # not a comment b", "x"]),
    ("a: some words\n  continued here\n\n  after blank\nb:\n- item one\n  more\n- 12\nc: 1\n",
     {"a": "some words continued here\nafter blank", "b": ["item one more", 12], "c": 1}),
    # tags
    ("a: !!str 123\nb: !!int \"7\"\nc: !!float 1\nd: !custom x\ne: !!str true\nf: !!null ''\n", {"a": "123", "b": 7, "c": 1.0, "d": "x", "e": "true", "f": None}),
    ("a: !!map {k: v}\nb: !!seq [1]\n", {"a": {"k": "v"}, "b": [1]}),
    # numbers as yaml.v2 resolves them (resolve.go:147-190): ParseInt with base 0 after stripping '_', then yamlStyleFloat
    ("- 017\n- 0o17\n- 0b11\n- -0b11\n- 0x1F\n- 1_000\n- +12\n- 08\n- 1e3\n- 0E0\n- .5\n- 5.\n- _1\n- _\n- 1e\n- 0x\n- 1-2\n",
     [15, 15, 3, -3, 31, 1000, 12, 8.0, 1000.0, 0.0, 0.5, 5.0, "_1", "_", "1e", "0x", "1-2"]),
    # timestamps (yaml.v2 resolve.go:224-258) arrive as the RFC 3339 string the reference's YAML -> JSON step makes of them
    ("- 2001-12-14\n- 2001-12-14T21:59:43.10-05:00\n- 2001-12-14t21:59:43.1Z\n- 2001-12-14 21:59:43.10\n- 2001-2-3\n- 2001-02-30\n"
     "- 2024-gpu\n- 2001-12-14T21:59:43\n- \"2001-12-14\"\n- 2001-12-14T1:2:3+00:00\n",
     ["2001-12-14T00:00:00Z", "2001-12-14T21:59:43.1-05:00", "2001-12-14T21:59:43.1Z", "2001-12-14T21:59:43.1Z", "2001-02-03T00:00:00Z",
      "2001-02-30", "2024-gpu", "2001-12-14T21:59:43", "2001-12-14", "2001-12-14T01:02:03Z"]),
    # mapping keys are resolved, then turned back into strings (sigs.k8s.io/yaml yaml.go:176-206)
    ("010: a\nyes: b\n\"no\": c\nplain: d\nm: {0x10: e, Off: f, \"010\": g}\n",
     {"8": "a", "true": "b", "no": "c", "plain": "d", "m": {"16": "e", "false": "f", "010": "g"}}),
    # directives, explicit start with a value, document end, a second document (ignored, like yaml.v2's Unmarshal)
    ("%YAML 1.1\n---\na: 1\n", {"a": 1}),
    ("--- \na: 1\n...\n", {"a": 1}),
    ("a: 1\n---\nb: 2\n", {"a": 1}),
    ("---\nversion: v1\n---\nversion: v2\n", {"version": "v1"}),
]


@pytest.mark.parametrize("text,want", HAND, ids=[str(i) for i in range(len(HAND))])
def test_hand_written_documents(text, want):
    status, got = product_parse(text)
    assert status == "ok", got
    assert same(got, want), (text, got)


@pytest.mark.parametrize("text,needle", [
    ("a: *nope\n", "unknown anchor"),
    ("~: x\n", "unsupported map key"),
    ("a: &x 1\nb:\n  <<: *x\n", "map merge requires"),
    ("a: !!binary aGk=\n", "unsupported YAML feature"),
    ("a: |x\n  t\n", "bad block scalar header"),
    ("a: \"never closed\n  at all\n", "unterminated"),
    ("a: plain scalar\n  nested: key\n", "mapping values are not allowed"),
    ("\tkey: v\n", "tabs"),
])
def test_refusals_are_loud(text, needle):
    status, got = product_parse(text)
    assert status == "error" and needle in got, got


def test_config_file_with_anchors_and_block_scalars_reaches_the_labeler(tmp_path):
    """End to end through the binary: a ConfigMap-style file using an anchor, an alias, a merge key and a second document."""
    import nodes as N
    from shimutil import run_cli, use_node
    cfg = os.path.join(str(tmp_path), "config.yaml")
    open(cfg, "w").write("%YAML 1.1\n---\nversion: v1\nflags: &f\n  migStrategy: mixed\n  gfd: {oneshot: true, noTimestamp: true}\n"
                         "sharing:\n  timeSlicing:\n    resources:\n    - &r\n      name: nvidia.com/gpu\n      replicas: 2\n"
                         "    - <<: *r\n      name: nvidia.com/mig-1g.23gb\n---\nversion: v9\n")
    fix = use_node(N.nodes()["c2_1xB200"], tmp_path)
    empty = os.path.join(str(tmp_path), "pci")
    os.makedirs(empty)
    out = os.path.join(str(tmp_path), "gfd")
    rc, _, se = run_cli(["--config-file", cfg, "-o", out], env={"GFD_FAKE_NVML": fix, "GFD_B200_PCI_DEVICES_ROOT": empty})
    assert rc == 0, se
    got = open(out).read()
    assert "nvidia.com/gpu.replicas=2\n" in got and "nvidia.com/gpu.product=NVIDIA-B200-SHARED\n" in got and "nvidia.com/mig.strategy=mixed\n" in got


@settings(max_examples=500, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(text=st.text(alphabet="ab01 \n\t-:&*!|>'\"[]{},#%<?.~+_\\", min_size=0, max_size=60))
def test_reader_never_crashes_or_hangs(text):
    """Arbitrary YAML-looking input: a value or a clean parse error, never a signal, never a hang (a config file is user input)."""
    p = subprocess.run([DUMP], input=text.encode(), capture_output=True, timeout=20)
    assert p.returncode in (0, 1), (text, p.returncode, p.stderr[-300:])
    if p.returncode == 1:
        assert p.stderr.decode().startswith("yaml: "), p.stderr
    else:
        json.loads(p.stdout.decode())
