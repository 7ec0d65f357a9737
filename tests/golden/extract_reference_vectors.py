#!/usr/bin/env python3
"""Extract the reference's own golden vectors into JSON fixtures.

TEST INFRASTRUCTURE.  Run in the build container (where /root/reference is
mounted); the GPU box never sees /root/reference, only the JSON this writes.

    python tests/golden/extract_reference_vectors.py [/root/reference]

What it reads (all paths under the reference root) and what it writes
(under tests/golden/):

  internal/lm/resource_test.go       TestGPUResourceLabeler, TestMigResourceLabeler
  internal/lm/mig-strategy_test.go   TestMigStrategyNoneLabels, TestMigStrategySingleLabels
  internal/lm/nvml_test.go           TestMigCapabilityLabeler
        -> lm_vectors.json  (the table-driven cases: inputs as a small device
           description, expected label map verbatim)
  cmd/gpu-feature-discovery/main_test.go   TestFailOnNVMLInitError table
        -> init_error_matrix.json
  internal/vgpu/pciutil.go:182-183   the two 256-byte PCI config-space
        known-answer blobs used by vgpu_test.go / pciutil_test.go
        -> vgpu_blobs.json (hex)
  tests/expected-output*.txt         the run()-level regex goldens
        -> expected_output_regexes.json ({file name: [regex lines]})

The Go test tables are parsed with a tiny recursive-descent parser for the
subset of Go expression syntax those tables use (composite literals, calls,
method chains, string/int/bool literals).  Nothing here is hand-typed: if the
reference's tables change, re-running this script changes the fixtures.
"""
import json
import os
import re
import sys

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))


# --------------------------------------------------------------------------
# A tokenizer + parser for the Go-expression subset used in the test tables.
# --------------------------------------------------------------------------
TOKEN_RE = re.compile(
    r"""
    (?P<ws>\s+|//[^\n]*)
  | (?P<str>"(?:\\.|[^"\\])*")
  | (?P<num>0x[0-9a-fA-F]+|\d+)
  | (?P<ident>[A-Za-z_][A-Za-z0-9_]*)
  | (?P<punct>\[\]|[{}()\[\],.:*&])
    """,
    re.X,
)


def tokenize(src):
    pos, out = 0, []
    while pos < len(src):
        m = TOKEN_RE.match(src, pos)
        if not m:
            raise SyntaxError("cannot tokenize at %r" % src[pos:pos + 40])
        pos = m.end()
        if m.lastgroup == "ws":
            continue
        out.append((m.lastgroup, m.group()))
    return out


class Parser:
    def __init__(self, toks):
        self.t, self.i = toks, 0

    def peek(self, k=0):
        return self.t[self.i + k] if self.i + k < len(self.t) else ("eof", "")

    def eat(self, val=None):
        kind, v = self.peek()
        if val is not None and v != val:
            raise SyntaxError("expected %r got %r at token %d" % (val, v, self.i))
        self.i += 1
        return kind, v

    # type := '[]' type | 'map' '[' type ']' type | ident ('.' ident)* | 'interface' '{' '}' | 'struct' ...
    def parse_type(self):
        kind, v = self.peek()
        if v == "[]":
            self.eat()
            return "[]" + self.parse_type()
        if v == "*":
            self.eat()
            return "*" + self.parse_type()
        if v == "map":
            self.eat()
            self.eat("[")
            k = self.parse_type()
            self.eat("]")
            return "map[%s]%s" % (k, self.parse_type())
        if v == "interface":
            self.eat()
            self.eat("{")
            self.eat("}")
            return "interface{}"
        name = self.eat()[1]
        while self.peek()[1] == "." and self.peek(1)[0] == "ident":
            self.eat()
            name += "." + self.eat()[1]
        return name

    def parse_expr(self):
        kind, v = self.peek()
        if kind == "str":
            self.eat()
            node = json.loads(v)
        elif kind == "num":
            self.eat()
            node = int(v, 0)
        elif v in ("true", "false"):
            self.eat()
            node = (v == "true")
        elif v == "nil":
            self.eat()
            node = None
        elif v == "&":
            self.eat()
            node = self.parse_expr()
        elif v == "{":
            node = {"lit": None, "elems": self.parse_lit_body()}
        else:
            typ = self.parse_type()
            if self.peek()[1] == "{":
                node = {"lit": typ, "elems": self.parse_lit_body()}
            elif self.peek()[1] == "(":
                node = {"call": typ, "args": self.parse_args()}
            else:
                node = {"ident": typ}
        # method chains / postfix calls
        while self.peek()[1] == "." and self.peek(1)[0] == "ident" and self.peek(2)[1] == "(":
            self.eat()
            name = self.eat()[1]
            node = {"method": name, "recv": node, "args": self.parse_args()}
        return node

    def parse_args(self):
        self.eat("(")
        args = []
        while self.peek()[1] != ")":
            args.append(self.parse_expr())
            if self.peek()[1] == ".":  # variadic '...'
                while self.peek()[1] == ".":
                    self.eat()
            if self.peek()[1] == ",":
                self.eat()
        self.eat(")")
        return args

    def parse_lit_body(self):
        self.eat("{")
        elems = []
        while self.peek()[1] != "}":
            e = self.parse_expr()
            if self.peek()[1] == ":":
                self.eat()
                val = self.parse_expr()
                key = e["ident"] if isinstance(e, dict) and "ident" in e else e
                elems.append({"key": key, "value": val})
            else:
                elems.append({"value": e})
            if self.peek()[1] == ",":
                self.eat()
        self.eat("}")
        return elems


def find_table(src, func_name):
    """Return the source text of the `testCases := []struct{...}{ ... }` value inside func_name."""
    start = src.index("func %s(" % func_name)
    m = re.compile(r"testCases\s*:=\s*\[\]struct\s*\{").search(src, start)
    # skip the struct type body
    i = m.end()
    depth = 1
    while depth:
        c = src[i]
        depth += (c == "{") - (c == "}")
        i += 1
    # now at the literal body '{'
    j = src.index("{", i)
    k, depth = j + 1, 1
    while depth:
        c = src[k]
        if c == '"':
            k += 1
            while src[k] != '"':
                k += 2 if src[k] == "\\" else 1
        depth += (c == "{") - (c == "}")
        k += 1
    return src[j:k]


def parse_table(src, func_name):
    body = find_table(src, func_name)
    p = Parser(tokenize(body))
    rows = p.parse_lit_body()
    out = []
    for r in rows:
        row = {}
        for e in r["value"]["elems"]:
            row[e["key"]] = e["value"]
        out.append(row)
    return out


# --------------------------------------------------------------------------
# Interpret the parsed AST nodes into plain fixture data.
# --------------------------------------------------------------------------
def lit_to_map(node):
    return {e["key"]: e["value"] for e in node["elems"]}


def conv_mig_device(node):
    assert node["call"] == "rt.NewMigDevice", node
    gi, ci, gb = node["args"][:3]
    attrs = {}
    for extra in node["args"][3:]:
        attrs.update(lit_to_map(extra))
    return {"gi": gi, "ci": ci, "gb": gb, "attributes": attrs}


def conv_device(node):
    """rt.NewFullGPU() | rt.NewMigEnabledDevice(migs...) | rt.NewDeviceMock(b).WithMigDevices(migs...)"""
    if "method" in node:
        assert node["method"] == "WithMigDevices"
        base = conv_device(node["recv"])
        base["migs"] = [conv_mig_device(a) for a in node["args"]]
        return base
    fn = node["call"]
    if fn == "rt.NewFullGPU":
        return {"mig_enabled": False, "migs": []}
    if fn == "rt.NewMigEnabledDevice":
        return {"mig_enabled": True, "migs": [conv_mig_device(a) for a in node["args"]]}
    if fn == "rt.NewDeviceMock":
        return {"mig_enabled": bool(node["args"][0]), "migs": []}
    raise ValueError(fn)


def conv_time_slicing(node):
    if node is None:
        return {"resources": []}
    m = lit_to_map(node)
    res = []
    for e in m.get("Resources", {"elems": []})["elems"]:
        r = lit_to_map(e["value"])
        res.append({"name": r["Name"], "rename": r.get("Rename", ""), "replicas": r["Replicas"]})
    return {"resources": res, "renameByDefault": bool(m.get("RenameByDefault", False))}


def conv_labels(node):
    if node is None:
        return None
    return lit_to_map(node)


def conv_rows(rows, kind):
    out = []
    for r in rows:
        o = {"description": r["description"]}
        if "devices" in r:
            o["devices"] = [conv_device(e["value"]) for e in r["devices"]["elems"]]
        elif kind in ("none", "single", "migcap"):
            o["devices"] = []
        if kind in ("gpu_resource", "mig_resource", "none"):
            o["timeSlicing"] = conv_time_slicing(r.get("timeSlicing"))
        if "count" in r or kind in ("gpu_resource", "mig_resource"):
            o["count"] = r.get("count", 0)
        if kind == "mig_resource":
            o["resourceName"] = r.get("resourceName", "")
        o["expectedError"] = bool(r.get("expectedError", False))
        o["expectedLabels"] = conv_labels(r.get("expectedLabels"))
        out.append(o)
    return out


def read(rel):
    with open(os.path.join(REF, rel)) as f:
        return f.read()


def main():
    vectors = {"_source": "generated by tests/golden/extract_reference_vectors.py from the reference's *_test.go tables"}

    src = read("internal/lm/resource_test.go")
    vectors["TestGPUResourceLabeler"] = {
        "cite": "internal/lm/resource_test.go:27-132",
        "device": {"mig_enabled": False, "migs": []},  # rt.NewFullGPU(), resource_test.go:28
        "cases": conv_rows(parse_table(src, "TestGPUResourceLabeler"), "gpu_resource"),
    }
    # resource_test.go:136-137: device := rt.NewMigDevice(1, 2, 300); rt.NewMigEnabledDevice(device)
    m = re.search(r"device := (rt\.NewMigDevice\([^)]*\))", src)
    migdev = conv_mig_device(Parser(tokenize(m.group(1))).parse_expr())
    vectors["TestMigResourceLabeler"] = {
        "cite": "internal/lm/resource_test.go:134-304",
        "device": migdev,
        "cases": conv_rows(parse_table(src, "TestMigResourceLabeler"), "mig_resource"),
    }

    src = read("internal/lm/mig-strategy_test.go")
    vectors["TestMigStrategyNoneLabels"] = {
        "cite": "internal/lm/mig-strategy_test.go:28-195",
        "strategy": "none",
        "cases": conv_rows(parse_table(src, "TestMigStrategyNoneLabels"), "none"),
    }
    vectors["TestMigStrategySingleLabels"] = {
        "cite": "internal/lm/mig-strategy_test.go:197-402",
        "strategy": "single",
        "cases": conv_rows(parse_table(src, "TestMigStrategySingleLabels"), "single"),
    }

    src = read("internal/lm/nvml_test.go")
    vectors["TestMigCapabilityLabeler"] = {
        "cite": "internal/lm/nvml_test.go:11-77",
        "cases": conv_rows(parse_table(src, "TestMigCapabilityLabeler"), "migcap"),
    }

    # The mock fixture constants the tables rely on (internal/resource/testing/resource-testing.go:31-134)
    rt = read("internal/resource/testing/resource-testing.go")
    vectors["mock"] = {
        "cite": "internal/resource/testing/resource-testing.go:31-134",
        "name": re.search(r'GetNameFunc: func\(\) \(string, error\) \{ return "([^"]+)"', rt).group(1),
        "total_memory_mb": int(re.search(r"return uint64\((\d+)\), nil", rt).group(1)),
        "cc_full": [int(x) for x in re.search(r"return (\d+), (\d+), nil\n\t\t\},", rt).groups()],
        "cc_mig_enabled": [0, 0],
        "driver_version": re.search(r'return "(\d+\.\d+)", nil', rt).group(1),
        "cuda_major": int(re.search(r"var major uint = (\d+)", rt).group(1)),
        "cuda_minor": int(re.search(r"var minor uint = (\d+)", rt).group(1)),
    }

    with open(os.path.join(OUT, "lm_vectors.json"), "w") as f:
        json.dump(vectors, f, indent=1, sort_keys=True)
        f.write("\n")

    # --- TestFailOnNVMLInitError matrix (main_test.go:273-380)
    src = read("cmd/gpu-feature-discovery/main_test.go")
    rows = parse_table(src, "TestFailOnNVMLInitError")
    matrix = []
    for r in rows:
        matrix.append({
            "description": r["description"],
            "errorOnInit": r["errorOnInit"] is not None,
            "failOnInitError": r["failOnInitError"],
            "migStrategy": r["migStrategy"],
            "expectError": r["expectError"],
        })
    with open(os.path.join(OUT, "init_error_matrix.json"), "w") as f:
        json.dump({"cite": "cmd/gpu-feature-discovery/main_test.go:273-380", "cases": matrix}, f, indent=1)
        f.write("\n")

    # --- vGPU known-answer blobs (internal/vgpu/pciutil.go:182-183) + expectations from the tests
    src = read("internal/vgpu/pciutil.go")
    blobs = {}
    for name in ("gpuPassThroughConfig", "vgpuConfig"):
        m = re.search(name + r"\s*=\s*\[\]byte\{([^}]*)\}", src)
        data = bytes(int(x, 16) for x in re.findall(r"0x([0-9a-fA-F]{2})", m.group(1)))
        blobs[name] = data.hex()
    tsrc = read("internal/vgpu/vgpu_test.go")
    expect = {
        "passthrough_cap_len": int(re.search(r"require\.Equal\(t, (\d+), len\(capability\), \"Vendor capability length for passthrough", tsrc).group(1)),
        "vgpu_cap_len": int(re.search(r"require\.Equal\(t, (\d+), len\(capability\), \"Vendor capability length for vgpu", tsrc).group(1)),
        "vgpu_cap_id": int(re.search(r"require\.Equal\(t, uint8\((\d+)\), GetByte\(capability, 0\)", tsrc).group(1)),
        "host_driver_version": re.search(r'require\.Equal\(t, "([^"]+)", info\.HostDriverVersion', tsrc).group(1),
        "host_driver_branch": re.search(r'require\.Equal\(t, "([^"]+)", info\.HostDriverBranch', tsrc).group(1),
        "config_len": int(re.search(r"len\(device\.pci\.Config\), (\d+)", tsrc).group(1)),
    }
    with open(os.path.join(OUT, "vgpu_blobs.json"), "w") as f:
        json.dump({
            "cite": "internal/vgpu/pciutil.go:180-204 (blobs); internal/vgpu/vgpu_test.go:36-74, pciutil_test.go:26-42 (expectations)",
            "blobs": blobs, "expect": expect}, f, indent=1)
        f.write("\n")

    # --- regex goldens (tests/expected-output*.txt): one JSON holding every file's regex lines
    goldens = {"cite": "tests/expected-output.txt, -mig-none.txt, -mig-single.txt, -mig-mixed.txt, -vgpu.txt (used by "
                       "cmd/gpu-feature-discovery/main_test.go:403-435 and tests/integration-tests.py:12-33)", "files": {}}
    for fn in sorted(os.listdir(os.path.join(REF, "tests"))):
        if fn.startswith("expected-output") and fn.endswith(".txt"):
            with open(os.path.join(REF, "tests", fn)) as f:
                goldens["files"][fn] = f.read().rstrip("\n").split("\n")
    with open(os.path.join(OUT, "expected_output_regexes.json"), "w") as f:
        json.dump(goldens, f, indent=1)
        f.write("\n")

    # --- the node labels the reference's README shows for a real A100 node (README.md:137-151): the only place the
    # reference documents VALUES that come out of the NVML boundary (two-part driver version -> empty rev, CUDA version
    # split, memory in MiB, family from compute capability, product from the device name)
    readme = read("README.md").split("\n")
    first = next(i for i, l in enumerate(readme) if l.strip() == "labels:")
    labels = {}
    for l in readme[first + 1:]:
        m = re.match(r'\s+(nvidia\.com/[\w.\-]+): "?([^"]*)"?\s*$', l)
        if not m:
            break
        labels[m.group(1)] = m.group(2)
    with open(os.path.join(OUT, "readme_node_example.json"), "w") as f:
        json.dump({"cite": "README.md:%d-%d" % (first + 2, first + 1 + len(labels)), "labels": labels}, f, indent=1, sort_keys=True)
        f.write("\n")

    # --- how the reference's own manifests start the binary: container args + env of every static manifest / job template
    manifests = {}
    rels = ["gpu-feature-discovery-daemonset.yaml", "gpu-feature-discovery-job.yaml.template"]
    rels += sorted("deployments/static/" + f for f in os.listdir(os.path.join(REF, "deployments", "static"))
                   if f.startswith("gpu-feature-discovery") and f.endswith((".yaml", ".template")))
    for rel in rels:
        src = read(rel)
        env = re.findall(r'- name: ([A-Z][A-Z0-9_]*)\s*\n\s*value: "?([^"\n]*)"?', src)
        args = []
        m = re.search(r"\n\s*args:\s*\n((?:\s*- .*\n)+)", src)
        if m:
            args = [a.strip()[2:].strip().strip('"') for a in m.group(1).strip("\n").split("\n")]
        mounts = re.findall(r'mountPath: "?([^"\n]+)"?', src)
        manifests[rel] = {"env": dict(env), "args": args, "mount_paths": mounts}
    with open(os.path.join(OUT, "manifests.json"), "w") as f:
        json.dump({"cite": "container spec of each manifest under the reference root", "manifests": manifests}, f, indent=1, sort_keys=True)
        f.write("\n")

    n = sum(len(v["cases"]) for v in vectors.values() if isinstance(v, dict) and "cases" in v)
    print("wrote %d lm cases, %d init-matrix rows, 2 vGPU blobs" % (n, len(matrix)))


if __name__ == "__main__":
    main()
