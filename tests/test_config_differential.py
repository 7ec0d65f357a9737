"""Config-file handling of the binary vs oracle/config_oracle.py on GENERATED documents (hypothesis).

A document is drawn as a Python object (right and wrong types in every position the Go code type-checks), rendered
as YAML by PyYAML in a randomly chosen style (block / flow / JSON, different indents, quoting as PyYAML decides) and
given to `gpu-feature-discovery --config-file`.  The config the binary logs ("Running with config:") must equal what
the oracle derives from the same object and command line, or both must reject the document.  This exercises the
labeler's own YAML reader (labeler/yaml_lite.cpp) against an independent YAML implementation's output."""
import json
import os
import re

import yaml
from hypothesis import HealthCheck, given, settings, strategies as st

import nodes as N
from oracle import config_oracle as C
from shimutil import run_cli, use_node


class _NoAlias(yaml.SafeDumper):
    def ignore_aliases(self, data):
        return True


TEXT = st.text(alphabet="abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789-_./: #'\"[]{},&*!|>%@`", min_size=0, max_size=12)
NAMES = st.sampled_from(["nvidia.com/gpu", "gpu", "nvidia.com/mig-1g.23gb", "mig-3g.90gb", "nvidia.com/GPU", "x" * 70, "a..b", "", "gpu.shared", "-gpu"])
SCALARS = st.one_of(st.none(), st.booleans(), st.integers(-5, 10 ** 12), st.sampled_from([0.5, 2.0, 60.0, 1.5e9]), TEXT)
DURATIONS = st.one_of(st.sampled_from(["60s", "90s", "1h2m3.5s", "1.5s", "250ms", "0", "10us", "soon", "1 s", "", "-5s", "1h", "100ns", "2m"]),
                      st.integers(0, 10 ** 11), st.sampled_from([1.5e9, 2.0]), st.none(), st.booleans())
DEVICES = st.one_of(st.sampled_from(["all", "some", 1, 3, 0, -1, 2.0, None, True]),
                    st.lists(st.sampled_from([0, 1, 7, "2", "0:1", "x", "GPU-b1028956-cfa2-0990-bf4a-5da9abb51763", "MIG-b1028956-cfa2-0990-bf4a-5da9abb51763",
                                              "MIG-GPU-b1028956-cfa2-0990-bf4a-5da9abb51763/3/0", -1, 1.5, None]), max_size=3))


def maybe(d, key, strat, p=0.6):
    return st.one_of(st.just({}), strat.map(lambda v: {key: v})) if p else strat.map(lambda v: {key: v})


@st.composite
def dict_of(draw, fields):
    out = {}
    for key, strat in fields.items():
        if draw(st.booleans()):
            k = key
            if draw(st.integers(0, 9)) == 0:
                k = key.upper() if draw(st.booleans()) else key[0].upper() + key[1:]
            out[k] = draw(strat)
    return out


REPLICATED = st.one_of(st.fixed_dictionaries({}, optional={"name": st.one_of(NAMES, st.integers(0, 3), st.none()), "rename": st.one_of(NAMES, st.none(), st.integers(0, 2)),
                                                             "devices": DEVICES, "replicas": st.one_of(st.integers(-1, 6), st.sampled_from([2.0, 2.5, "2", None, True]))}),
                       st.sampled_from([None, 3, "x"]))
TIMESLICING = st.one_of(st.fixed_dictionaries({}, optional={"renameByDefault": st.one_of(st.booleans(), st.sampled_from(["true", 1, None])),
                                                             "failRequestsGreaterThanOne": st.one_of(st.booleans(), st.sampled_from(["no", 0])),
                                                             "resources": st.one_of(st.lists(REPLICATED, max_size=3), st.sampled_from([None, "x", {}]))}),
                        st.sampled_from([None, [], "x"]))
RESOURCE = st.one_of(st.fixed_dictionaries({}, optional={"pattern": st.one_of(TEXT, st.integers(0, 3), st.none()), "name": st.one_of(NAMES, st.none(), st.booleans())}),
                     st.sampled_from([None, "x"]))

GFD = dict_of({"oneshot": st.one_of(st.booleans(), st.sampled_from(["true", 1, None])), "noTimestamp": st.one_of(st.booleans(), st.sampled_from(["x", None])),
               "sleepInterval": DURATIONS, "outputFile": SCALARS, "machineTypeFile": SCALARS})
PLUGIN = dict_of({"passDeviceSpecs": st.one_of(st.booleans(), st.just("yes")), "deviceListStrategy": SCALARS, "deviceIDStrategy": SCALARS})
FLAGS = dict_of({"migStrategy": SCALARS, "failOnInitError": st.one_of(st.booleans(), st.sampled_from(["true", 0, None])), "nvidiaDriverRoot": SCALARS,
                 "gdsEnabled": st.one_of(st.booleans(), st.none(), st.just(1)), "mofedEnabled": st.one_of(st.booleans(), st.none()),
                 "plugin": st.one_of(PLUGIN, st.sampled_from([None, 1])), "gfd": st.one_of(GFD, st.sampled_from([None, "x", []]))})
DOC = st.one_of(dict_of({"version": st.one_of(st.sampled_from(["v1", "v2", "", None, 1, True]), TEXT),
                         "flags": st.one_of(FLAGS, st.sampled_from([None, 7])),
                         "resources": st.one_of(dict_of({"gpus": st.one_of(st.lists(RESOURCE, max_size=2), st.sampled_from([None, 3])),
                                                         "mig": st.one_of(st.lists(RESOURCE, max_size=2), st.none())}), st.sampled_from([None, "r"])),
                         "sharing": st.one_of(dict_of({"timeSlicing": TIMESLICING}), st.sampled_from([None, 0]))}),
                st.sampled_from([None, [], "text", 5]))
STYLE = st.sampled_from(["block", "block4", "flow", "json", "json_indent"])
CLI = st.fixed_dictionaries({}, optional={"migStrategy": st.sampled_from(["none", "single", "mixed", ""]), "failOnInitError": st.booleans(),
                                          "noTimestamp": st.booleans(), "sleepInterval": st.sampled_from([10 ** 9, 90 * 10 ** 9, 1500 * 10 ** 6]),
                                          "machineTypeFile": st.sampled_from(["", "/dev/null"])})


def render(doc, style):
    if style == "json":
        return json.dumps(doc)
    if style == "json_indent":
        return json.dumps(doc, indent=3)
    kw = {"block": dict(default_flow_style=False), "block4": dict(default_flow_style=False, indent=4), "flow": dict(default_flow_style=True)}[style]
    return yaml.dump(doc, Dumper=_NoAlias, width=100000, **kw)


def yaml_types_survive(doc, text):
    """Only keep examples whose rendering decodes (with PyYAML) back to the same object: the product is compared
    on what the document SAYS, and that must not depend on PyYAML-vs-yaml.v2 corner cases."""
    try:
        return yaml.safe_load(text) == doc
    except yaml.YAMLError:
        return False


@settings(max_examples=600, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow, HealthCheck.filter_too_much])
@given(doc=DOC, style=STYLE, cli=CLI)
def test_config_file_matches_oracle(doc, style, cli, tmp_path_factory):
    text = render(doc, style)
    if not yaml_types_survive(doc, text):
        return
    tmp = str(tmp_path_factory.mktemp("cfg"))
    cfg = os.path.join(tmp, "config.yaml")
    open(cfg, "w").write(text)
    out = os.path.join(tmp, "gfd")
    fix = use_node(N.nodes()["zero_gpus"], tmp)
    empty = os.path.join(tmp, "pci")
    os.makedirs(empty)
    args = ["--oneshot", "-o", out, "--config-file", cfg]
    cli_set = {"oneshot": True, "outputFile": out}
    if "migStrategy" in cli:
        args.append("--mig-strategy=" + cli["migStrategy"])
    if "failOnInitError" in cli:
        args.append("--fail-on-init-error=%s" % str(cli["failOnInitError"]).lower())
    if "noTimestamp" in cli:
        args.append("--no-timestamp=%s" % str(cli["noTimestamp"]).lower())
    if "sleepInterval" in cli:
        args += ["--sleep-interval", C.duration_string(cli["sleepInterval"])]
    if "machineTypeFile" in cli:
        args.append("--machine-type-file=" + cli["machineTypeFile"])
    cli_set.update(cli)
    try:
        want = C.load_config(True, doc, cli_set)
    except C.ConfigError:
        want = None
    rc, _, se = run_cli(args, env={"GFD_FAKE_NVML": fix, "GFD_B200_PCI_DEVICES_ROOT": empty})
    m = re.search(r"Running with config:\n(\{.*?\n\})\n", se, re.S)
    if want is None:
        assert rc == 1 and m is None and "unable to parse config file" in se, (text, se[-600:])
    else:
        assert m is not None, (text, se[-600:])
        assert json.loads(m.group(1)) == want, text
        assert rc == 0, (text, se[-600:])
