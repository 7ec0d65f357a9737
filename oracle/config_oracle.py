"""Config oracle -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Restates, over plain Python objects (what YAML/JSON decoding yields), how the reference turns a config document plus
command-line flags into the `*spec.Config` it logs with json.MarshalIndent:

    spec.NewConfig / parseConfigFrom              V/spec/config.go:43-99   (V/spec = vendor/github.com/NVIDIA/k8s-device-plugin/api/config/v1)
    Flags.UpdateFromCLIFlags / updateFromCLIFlag  V/spec/flags.go:29-40,75-121
    TimeSlicing / ReplicatedResource / ReplicatedDevices .UnmarshalJSON   V/spec/replicas.go:121-281
    Resource / ResourceName .UnmarshalJSON, NewResourceName               V/spec/resources.go:48-64,120-165
    Duration.UnmarshalJSON / MarshalJSON                                  V/spec/duration.go:28-51
    loadConfig, disableResourceRenamingInConfig   cmd/gpu-feature-discovery/main.go:96-107,236-270
    sigs.k8s.io/yaml: YAML scalars are coerced to strings when the Go target is a plain string

PARITY UNPINNED: the reference has no test for any of this; the restatement is checked only against hand-read
expectations (tests/test_cli_flags_and_lifecycle.py) and, differentially, against the product on generated documents
(tests/test_config_differential.py) -- which pins the product's YAML reader and its Go-JSON typing rules to THIS
reading of the Go code, not to the Go code itself.
"""
import math
import re


class ConfigError(Exception):
    pass


_DNS_LABEL = r"[a-z0-9]([-a-z0-9]*[a-z0-9])?"
_DNS_SUBDOMAIN = re.compile(r"^%s(\.%s)*$" % (_DNS_LABEL, _DNS_LABEL))
_UUID = re.compile(r"^[0-9a-fA-F]{8}-[0-9a-fA-F]{4}-[0-9a-fA-F]{4}-[0-9a-fA-F]{4}-[0-9a-fA-F]{12}$")
_DUR = re.compile(r"^[-+]?(([0-9]+(\.[0-9]*)?|\.[0-9]+)(ns|us|µs|μs|ms|s|m|h))+$")
_DUR_PART = re.compile(r"([0-9]*\.?[0-9]*)(ns|us|µs|μs|ms|s|m|h)")
_UNIT = {"ns": 1, "us": 1000, "µs": 1000, "μs": 1000, "ms": 10 ** 6, "s": 10 ** 9, "m": 60 * 10 ** 9, "h": 3600 * 10 ** 9}


def parse_duration(s):
    """time.ParseDuration (ns) for the well-formed inputs the tests generate; ConfigError otherwise."""
    if s in ("0", "+0", "-0"):
        return 0
    if not _DUR.match(s):
        raise ConfigError("time: invalid duration %r" % s)
    neg = s.startswith("-")
    total = 0
    for num, unit in _DUR_PART.findall(s.lstrip("+-")):
        ip, _, fp = num.partition(".")
        v = int(ip or "0") * _UNIT[unit]
        if fp:
            v += int(float(int(fp)) * (float(_UNIT[unit]) / (10.0 ** len(fp))))
        total += v
    return -total if neg else total


def duration_string(d):
    """time.Duration.String()"""
    if d == 0:
        return "0s"
    neg, u = d < 0, abs(d)
    if u < 10 ** 9:
        if u < 1000:
            out = "%dns" % u
        elif u < 10 ** 6:
            out = _frac(u, 3) + "µs"
        else:
            out = _frac(u, 6) + "ms"
    else:
        secs_frac = _frac(u % (60 * 10 ** 9), 9)
        out = secs_frac + "s"
        mins = u // (60 * 10 ** 9)
        if mins > 0:
            out = "%dm" % (mins % 60) + out
            if mins // 60 > 0:
                out = "%dh" % (mins // 60) + out
    return ("-" if neg else "") + out


def _frac(v, prec):
    ip, fp = divmod(v, 10 ** prec)
    f = ("%0*d" % (prec, fp)).rstrip("0")
    return "%d" % ip + ("." + f if f else "")


def _kind(v):
    if v is None:
        return "null"
    if isinstance(v, bool):
        return "bool"
    if isinstance(v, (int, float)):
        return "number"
    if isinstance(v, str):
        return "string"
    if isinstance(v, list):
        return "array"
    return "object"


def _fmt_float32(f):
    for prec in range(1, 10):
        s = "%.*g" % (prec, f)
        import struct
        if struct.unpack("f", struct.pack("f", float(s)))[0] == struct.unpack("f", struct.pack("f", f))[0]:
            return s
    return "%.9g" % f


def _plain_string(v, field):
    """Go string target behind sigs.k8s.io/yaml: scalars become strings; null leaves the field unset (None)."""
    if v is None:
        return None
    if isinstance(v, bool):
        return "true" if v else "false"
    if isinstance(v, int):
        return str(v)
    if isinstance(v, float):
        return _fmt_float32(v)
    if isinstance(v, str):
        return v
    raise ConfigError("cannot unmarshal %s into %s of type string" % (_kind(v), field))


def _json_string(v, field):
    """a type with its own UnmarshalJSON reading a JSON string: no YAML coercion"""
    if v is None:
        return None
    if not isinstance(v, str):
        raise ConfigError("cannot unmarshal %s into Go value of type string (%s)" % (_kind(v), field))
    return v


def _bool(v, field):
    if v is None:
        return None
    if not isinstance(v, bool):
        raise ConfigError("cannot unmarshal %s into %s of type bool" % (_kind(v), field))
    return v


def _get_fold(obj, key):
    """encoding/json field matching: exact key first, else ASCII case-insensitive."""
    if key in obj:
        return True, obj[key]
    for k, v in obj.items():
        if isinstance(k, str) and k.lower() == key.lower():
            return True, v
    return False, None


def new_resource_name(n):
    if not n.startswith("nvidia.com/"):
        n = "nvidia.com/" + n
    if len(n) > 63:
        raise ConfigError("fully-qualified resource name must be 63 characters or less: %s" % n)
    name = n.split("/", 1)[1]
    if len(name) > 253 or not _DNS_SUBDOMAIN.match(name):
        raise ConfigError("incorrect format for resource name '%s'" % n)
    return n


def _is_uint(s):
    return bool(re.match(r"^\+?[0-9]{1,19}$", s))


def _is_gpu_uuid(s):
    return s.startswith("GPU-") and bool(_UUID.match(s[4:]))


def _is_mig_uuid(s):
    if not s.startswith("MIG-"):
        return False
    suf = s[4:]
    if _UUID.match(suf):
        return True
    parts = suf.split("/", 2)
    return len(parts) == 3 and _is_gpu_uuid(parts[0]) and _is_uint(parts[1]) and _is_uint(parts[2])


def _devices(v):
    if isinstance(v, str):
        if v != "all":
            raise ConfigError("devices set as '%s' but the only valid string input is 'all'" % v)
        return "all"
    if isinstance(v, bool) or v is None:
        raise ConfigError("unrecognized type for devices spec")
    if isinstance(v, int) or (isinstance(v, float) and v == math.floor(v)):
        if int(v) <= 0:
            raise ConfigError("a count of devices must be > 0")
        return int(v)
    if isinstance(v, list):
        out = []
        for e in v:
            if isinstance(e, int) and not isinstance(e, bool) and e >= 0:
                out.append(str(e))
            elif isinstance(e, str) and (_is_uint(e) or (":" in e and all(_is_uint(p) for p in e.split(":", 1))) or _is_gpu_uuid(e) or _is_mig_uuid(e)):
                out.append(e)
            else:
                raise ConfigError("unsupported type for device in devices list")
        return out
    raise ConfigError("unrecognized type for devices spec")


def _replicated_resource(v):
    if v is not None and not isinstance(v, dict):
        raise ConfigError("cannot unmarshal %s into map" % _kind(v))
    v = v or {}
    if "name" not in v:
        raise ConfigError("no resource name specified")
    r = {"name": "", "rename": "", "devices": "all", "replicas": 0}
    name = _json_string(v["name"], "name")
    if name is not None:
        r["name"] = new_resource_name(name)
    if "devices" in v:
        r["devices"] = _devices(v["devices"])
    if "replicas" not in v:
        raise ConfigError("no replicas specified")
    rep = v["replicas"]
    if rep is not None:
        if isinstance(rep, bool) or not isinstance(rep, (int, float)) or (isinstance(rep, float) and rep != math.floor(rep)):
            raise ConfigError("cannot unmarshal %s into Go value of type int" % _kind(rep))
        r["replicas"] = int(rep)
    if r["replicas"] < 2:
        raise ConfigError("number of replicas must be >= 2")
    if "rename" in v:
        rn = _json_string(v["rename"], "rename")
        if rn is not None:
            r["rename"] = new_resource_name(rn)
    return r


def _time_slicing(v):
    if v is not None and not isinstance(v, dict):
        raise ConfigError("cannot unmarshal %s into map" % _kind(v))
    v = v or {}
    t = {"renameByDefault": False, "failRequestsGreaterThanOne": False, "resources": []}
    for k in ("renameByDefault", "failRequestsGreaterThanOne"):
        if k in v:
            b = _bool(v[k], k)
            if b is not None:
                t[k] = b
    if "resources" not in v:
        raise ConfigError("no resources specified")
    res = v["resources"]
    if res is not None and not isinstance(res, list):
        raise ConfigError("cannot unmarshal %s into []ReplicatedResource" % _kind(res))
    t["resources"] = [_replicated_resource(e) for e in (res or [])]
    if not t["resources"]:
        raise ConfigError("no resources specified")
    for r in t["resources"]:
        if t["renameByDefault"] and r["rename"] == "":
            r["rename"] = r["name"] + ".shared"
    return t


def _resource(v):
    if v is not None and not isinstance(v, dict):
        raise ConfigError("cannot unmarshal %s into map" % _kind(v))
    v = v or {}
    if "pattern" not in v:
        raise ConfigError("resources must have a 'pattern' field set")
    if "name" not in v:
        raise ConfigError("resources must have a 'name' field set")
    _json_string(v["pattern"], "pattern")
    n = _json_string(v["name"], "name")
    if n is not None:
        new_resource_name(n)
    return True


def parse_config(doc):
    """parseConfigFrom over the decoded document -> internal dict with None for unset pointer fields."""
    c = {"version": "", "migStrategy": None, "failOnInitError": None, "nvidiaDriverRoot": None, "gdsEnabled": None, "mofedEnabled": None,
         "oneshot": None, "noTimestamp": None, "sleepInterval": None, "outputFile": None, "machineTypeFile": None,
         "timeSlicing": {"renameByDefault": False, "failRequestsGreaterThanOne": False, "resources": []}, "n_resources": 0}
    if doc is None:
        c["version"] = "v1"
        return c
    if not isinstance(doc, dict):
        raise ConfigError("cannot unmarshal %s into Go value of type v1.Config" % _kind(doc))
    ok, v = _get_fold(doc, "version")
    if ok:
        c["version"] = _plain_string(v, "version") or ""
    ok, f = _get_fold(doc, "flags")
    if ok and f is not None:
        if not isinstance(f, dict):
            raise ConfigError("cannot unmarshal %s into flags" % _kind(f))
        for key in ("migStrategy", "nvidiaDriverRoot"):
            ok2, v = _get_fold(f, key)
            if ok2:
                c[key] = _plain_string(v, key)
        for key in ("failOnInitError", "gdsEnabled", "mofedEnabled"):
            ok2, v = _get_fold(f, key)
            if ok2:
                c[key] = _bool(v, key)
        ok2, p = _get_fold(f, "plugin")
        if ok2 and p is not None:
            if not isinstance(p, dict):
                raise ConfigError("cannot unmarshal %s into plugin" % _kind(p))
            ok3, v = _get_fold(p, "passDeviceSpecs")
            if ok3:
                _bool(v, "passDeviceSpecs")
            for key in ("deviceListStrategy", "deviceIDStrategy"):
                ok3, v = _get_fold(p, key)
                if ok3:
                    _plain_string(v, key)
        ok2, g = _get_fold(f, "gfd")
        if ok2 and g is not None:
            if not isinstance(g, dict):
                raise ConfigError("cannot unmarshal %s into gfd" % _kind(g))
            for key in ("oneshot", "noTimestamp"):
                ok3, v = _get_fold(g, key)
                if ok3:
                    c[key] = _bool(v, key)
            ok3, v = _get_fold(g, "sleepInterval")
            if ok3:
                if isinstance(v, bool) or v is None or isinstance(v, (list, dict)):
                    raise ConfigError("invalid duration")
                if isinstance(v, (int, float)):
                    c["sleepInterval"] = int(v)  # Duration(time.Duration(float64)): truncation toward zero
                else:
                    c["sleepInterval"] = parse_duration(v)
            for key in ("outputFile", "machineTypeFile"):
                ok3, v = _get_fold(g, key)
                if ok3:
                    c[key] = _plain_string(v, key)
    ok, r = _get_fold(doc, "resources")
    if ok and r is not None:
        if not isinstance(r, dict):
            raise ConfigError("cannot unmarshal %s into resources" % _kind(r))
        for key in ("gpus", "mig"):
            ok2, lst = _get_fold(r, key)
            if ok2 and lst is not None:
                if not isinstance(lst, list):
                    raise ConfigError("cannot unmarshal %s into []Resource" % _kind(lst))
                for e in lst:
                    _resource(e)
                c["n_resources"] += len(lst)
    ok, s = _get_fold(doc, "sharing")
    if ok and s is not None:
        if not isinstance(s, dict):
            raise ConfigError("cannot unmarshal %s into sharing" % _kind(s))
        ok2, ts = _get_fold(s, "timeSlicing")
        if ok2:
            c["timeSlicing"] = _time_slicing(ts)
    if c["version"] == "":
        c["version"] = "v1"
    if c["version"] != "v1":
        raise ConfigError("unknown version: %s" % c["version"])
    return c


DEFAULTS = {"migStrategy": "none", "failOnInitError": True, "oneshot": False, "noTimestamp": False, "sleepInterval": 60 * 10 ** 9,
            "outputFile": "/etc/kubernetes/node-feature-discovery/features.d/gfd", "machineTypeFile": "/sys/class/dmi/id/product_name"}


def load_config(doc_present, doc, cli):
    """loadConfig + disableResourceRenamingInConfig.  `cli`: {flag key: value} for flags SET on the command line or
    through the environment (cli.Context.IsSet).  Returns the dict json.MarshalIndent would print."""
    c = parse_config(doc) if doc_present else parse_config(None)
    for key, default in DEFAULTS.items():
        if key in cli:
            c[key] = cli[key]
        elif c[key] is None:
            c[key] = default
    ts = c["timeSlicing"]
    for r in ts["resources"]:
        if not ts["renameByDefault"] and r["rename"] != "":
            r["rename"] = ""
        if ts["renameByDefault"] and r["rename"] != r["name"] + ".shared":
            r["rename"] = r["name"] + ".shared"
        r["devices"] = "all"
    flags = {"migStrategy": c["migStrategy"], "failOnInitError": c["failOnInitError"]}
    if c["nvidiaDriverRoot"] is not None:
        flags["nvidiaDriverRoot"] = c["nvidiaDriverRoot"]
    flags.update({"gdsEnabled": c["gdsEnabled"], "mofedEnabled": c["mofedEnabled"],
                  "gfd": {"oneshot": c["oneshot"], "noTimestamp": c["noTimestamp"], "sleepInterval": duration_string(c["sleepInterval"]),
                          "outputFile": c["outputFile"], "machineTypeFile": c["machineTypeFile"]}})
    tsj = {}
    if ts["renameByDefault"]:
        tsj["renameByDefault"] = True
    if ts["failRequestsGreaterThanOne"]:
        tsj["failRequestsGreaterThanOne"] = True
    if ts["resources"]:
        tsj["resources"] = []
        for r in ts["resources"]:
            e = {"name": r["name"]}
            if r["rename"]:
                e["rename"] = r["rename"]
            e.update({"devices": "all", "replicas": r["replicas"]})
            tsj["resources"].append(e)
    return {"version": c["version"], "flags": flags, "resources": {"gpus": None}, "sharing": {"timeSlicing": tsj}}
