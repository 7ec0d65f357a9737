"""Drop-in behaviour of the binary around the pass: flag grammar (urfave/cli v2 over Go's flag package), env-var
precedence, config dump, daemon loop, signals, output-file life-cycle (cmd/gpu-feature-discovery/main.go:22-232).
CPU only, fake NVML."""
import json
import os
import re
import signal
import subprocess
import time

import pytest

import nodes as N
from shimutil import B, binary_path, run_cli, use_node

NODES = N.nodes()


def base_env(tmp_path, node="c2_1xB200"):
    fix = use_node(NODES[node], tmp_path)
    empty = os.path.join(str(tmp_path), "pci-empty")
    os.makedirs(empty, exist_ok=True)
    return {"GFD_FAKE_NVML": fix, "GFD_B200_PCI_DEVICES_ROOT": empty}


def dumped_config(stderr):
    m = re.search(r"Running with config:\n(\{.*?\n\})\n", stderr, re.S)
    assert m, stderr
    return json.loads(m.group(1))


def test_config1_cpu_only_host(tmp_path):
    """BASELINE config #1: no NVML -> null manager -> 0-byte file, exit 0 (factory.go:42-74, nvml.go:40-42, main.go:175-187)."""
    out = os.path.join(str(tmp_path), "gfd")
    rc, so, se = run_cli(["--oneshot", "--no-timestamp", "--fail-on-init-error=false", "-o", out], fake=False,
                         env={"GFD_B200_PCI_DEVICES_ROOT": "/sys/bus/pci/devices"})
    assert rc == 0, se
    assert os.path.getsize(out) == 0 and oct(os.stat(out).st_mode & 0o777) == "0o644"
    assert "Detected non-NVML platform: could not load NVML" in se
    assert "WARNING: No valid resources detected; using empty manager." in se
    assert "Warning: no labels generated from any source" in se
    assert os.path.isdir(os.path.join(str(tmp_path), "gfd-tmp"))


def test_bogus_strategy_succeeds_without_devices(tmp_path):
    """main_test.go:308-319: the strategy is only validated once devices exist."""
    out = os.path.join(str(tmp_path), "gfd")
    rc, _, se = run_cli(["--oneshot", "--mig-strategy=bogus", "-o", out], fake=False, env={"GFD_B200_PCI_DEVICES_ROOT": "/sys/bus/pci/devices"})
    assert rc == 0, se
    assert re.fullmatch(r"nvidia\.com/gfd\.timestamp=[0-9]{10}\n", open(out).read())


def test_init_error_matrix_through_the_binary(tmp_path):
    """TestFailOnNVMLInitError (main_test.go:273-380) end to end: nvmlInit fails in the fake."""
    cases = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "init_error_matrix.json")))["cases"]
    for c in cases:
        env = base_env(tmp_path, "err_init" if c["errorOnInit"] else "c2_1xB200")
        out = os.path.join(str(tmp_path), "gfd-init")
        if os.path.exists(out):
            os.remove(out)
        rc, _, se = run_cli(["--oneshot", "--mig-strategy=" + c["migStrategy"], "--fail-on-init-error=%s" % str(c["failOnInitError"]).lower(),
                             "-o", out], env=env)
        assert (rc != 0) == c["expectError"], (c, se)
        if c["errorOnInit"] and not c["failOnInitError"]:
            assert "Warning: failed to initialize resource manager: ERROR_UNKNOWN" in se
            assert re.fullmatch(r"nvidia\.com/gfd\.timestamp=[0-9]{10}\n", open(out).read())


def test_output_to_stdout_when_path_is_empty(tmp_path):
    rc, so, se = run_cli(["--oneshot", "--no-timestamp", "--output-file="], env=base_env(tmp_path))
    assert rc == 0 and "nvidia.com/gpu.product=NVIDIA-B200\n" in so and so.count("\n") == 14


@pytest.mark.parametrize("args", [["-o", "X"], ["--output", "X"], ["--output-file", "X"], ["-output-file=X"], ["--o=X"]])
def test_output_flag_aliases(args, tmp_path):
    out = os.path.join(str(tmp_path), "f")
    rc, _, se = run_cli(["--oneshot", "--no-timestamp"] + [a.replace("X", out) for a in args], env=base_env(tmp_path))
    assert rc == 0 and os.path.getsize(out) > 0, se
    assert dumped_config(se)["flags"]["gfd"]["outputFile"] == out


def test_two_spellings_of_one_flag_is_an_error(tmp_path):
    rc, so, se = run_cli(["--oneshot", "-o", "/tmp/a", "--output", "/tmp/b"], env=base_env(tmp_path))
    assert rc == 1 and "Cannot use two forms of the same flag: output o" in se + so


@pytest.mark.parametrize("args,msg", [
    (["--nope"], "flag provided but not defined: -nope"),
    (["--oneshot=maybe"], 'invalid boolean value "maybe" for -oneshot: parse error'),
    (["--sleep-interval"], "flag needs an argument: -sleep-interval"),
    (["--sleep-interval=fast"], 'invalid value "fast" for flag -sleep-interval: parse error'),
    (["---x"], "bad flag syntax: ---x"),
])
def test_usage_errors(args, msg, tmp_path):
    rc, so, se = run_cli(args, env=base_env(tmp_path))
    assert rc == 1 and ("Error: " + msg) in se and ("Incorrect Usage. " + msg) in so and "GLOBAL OPTIONS:" in so


def test_help_and_version(tmp_path):
    for a in (["--help"], ["-h"], ["help"], ["h"]):
        rc, so, _ = run_cli(a, env=base_env(tmp_path))
        assert rc == 0 and "GPU Feature Discovery - generate labels for NVIDIA devices" in so and "--mig-strategy value" in so
    for a in (["--version"], ["-v"]):
        rc, so, _ = run_cli(a, env=base_env(tmp_path))
        assert rc == 0 and so.startswith("GPU Feature Discovery version ")


def test_bool_flag_does_not_take_next_argument(tmp_path):
    """Go's flag package: `--oneshot false` leaves oneshot=true and `false` positional (which also stops flag parsing)."""
    out = os.path.join(str(tmp_path), "f")
    rc, _, se = run_cli(["-o", out, "--no-timestamp", "--oneshot", "false", "--mig-strategy=single"], env=base_env(tmp_path))
    cfg = dumped_config(se)
    assert rc == 0 and cfg["flags"]["gfd"]["oneshot"] is True and cfg["flags"]["migStrategy"] == "none"


def test_env_vars_and_precedence(tmp_path):
    out = os.path.join(str(tmp_path), "f")
    env = base_env(tmp_path)
    env.update({"GFD_ONESHOT": "TRUE", "GFD_NO_TIMESTAMP": "1", "GFD_OUTPUT_FILE": out, "GFD_SLEEP_INTERVAL": "10s",
                "MIG_STRATEGY": "mixed", "GFD_FAIL_ON_INIT_ERROR": "f"})
    rc, _, se = run_cli([], env=env)
    cfg = dumped_config(se)
    assert rc == 0, se
    assert cfg["flags"] == {"migStrategy": "mixed", "failOnInitError": False, "gdsEnabled": None, "mofedEnabled": None,
                            "gfd": {"oneshot": True, "noTimestamp": True, "sleepInterval": "10s", "outputFile": out,
                                    "machineTypeFile": "/sys/class/dmi/id/product_name"}}
    # the code (not the README) decides: a flag on the command line beats the environment (V/cli/flag_bool.go:44-67)
    env["GFD_MIG_STRATEGY"] = "single"
    rc, _, se = run_cli(["--mig-strategy=none", "--sleep-interval", "1h2m3.5s"], env=env)
    cfg = dumped_config(se)
    assert cfg["flags"]["migStrategy"] == "none" and cfg["flags"]["gfd"]["sleepInterval"] == "1h2m3.5s"
    # first listed env var wins even when set to the empty string (string flags), empty bools are ignored
    env["GFD_MIG_STRATEGY"] = ""
    env["GFD_ONESHOT"] = ""
    rc, _, se = run_cli(["--oneshot"], env=env)
    assert dumped_config(se)["flags"]["migStrategy"] == ""


def test_bad_env_bool_is_fatal(tmp_path):
    env = base_env(tmp_path)
    env["GFD_ONESHOT"] = "yes"
    rc, so, se = run_cli([], env=env)
    assert rc == 1 and 'could not parse "yes" as bool value from environment variable "GFD_ONESHOT" for flag oneshot' in se


def test_config_file_precedence_and_dump(tmp_path):
    out = os.path.join(str(tmp_path), "f")
    cfgfile = os.path.join(str(tmp_path), "c.yaml")
    open(cfgfile, "w").write("""
version: v1
flags:
  migStrategy: "mixed"     # overridden by the CLI below
  failOnInitError: false
  nvidiaDriverRoot: "/run/nvidia/driver"
  gdsEnabled: true
  plugin:
    passDeviceSpecs: false
    deviceListStrategy: "envvar"
    deviceIDStrategy: "uuid"
  gfd:
    oneshot: true
    noTimestamp: true
    outputFile: %s
    sleepInterval: 90s
resources:
  gpus:
  - pattern: "*B200*"
    name: b200
sharing:
  timeSlicing:
    failRequestsGreaterThanOne: true
    resources:
    - name: nvidia.com/gpu
      rename: gpu-x
      devices: ["GPU-b1028956-cfa2-0990-bf4a-5da9abb51763", "0:1", 3]
      replicas: 4
""" % out)
    rc, _, se = run_cli(["--config-file", cfgfile, "--mig-strategy=none"], env=base_env(tmp_path))
    assert rc == 0, se
    assert dumped_config(se) == {
        "version": "v1",
        "flags": {"migStrategy": "none", "failOnInitError": False, "nvidiaDriverRoot": "/run/nvidia/driver", "gdsEnabled": True,
                  "mofedEnabled": None, "gfd": {"oneshot": True, "noTimestamp": True, "sleepInterval": "1m30s", "outputFile": out,
                                                  "machineTypeFile": "/sys/class/dmi/id/product_name"}},
        "resources": {"gpus": None},
        "sharing": {"timeSlicing": {"failRequestsGreaterThanOne": True,
                                    "resources": [{"name": "nvidia.com/gpu", "devices": "all", "replicas": 4}]}}}
    for needle in ("Customizing the 'resources' field is not yet supported", "Setting the 'rename' field in sharing.timeSlicing.resources",
                   "Customizing the 'devices' field in sharing.timeSlicing.resources"):
        assert needle in se
    text = open(out).read()
    assert "nvidia.com/gpu.replicas=4\n" in text and "nvidia.com/gpu.product=NVIDIA-B200-SHARED\n" in text


@pytest.mark.parametrize("doc,msg", [
    ("version: v2\n", "unknown version: v2"),
    ("sharing:\n  timeSlicing:\n    resources: []\n", "no resources specified"),
    ("sharing:\n  timeSlicing:\n    renameByDefault: true\n", "no resources specified"),
    ("sharing:\n  timeSlicing:\n    resources:\n    - name: nvidia.com/gpu\n      replicas: 1\n", "number of replicas must be >= 2"),
    ("sharing:\n  timeSlicing:\n    resources:\n    - name: nvidia.com/gpu\n", "no replicas specified"),
    ("sharing:\n  timeSlicing:\n    resources:\n    - replicas: 2\n", "no resource name specified"),
    ("sharing:\n  timeSlicing:\n    resources:\n    - name: nvidia.com/GPU\n      replicas: 2\n", "incorrect format for resource name"),
    ("sharing:\n  timeSlicing:\n    resources:\n    - name: " + "x" * 70 + "\n      replicas: 2\n", "must be 63 characters or less"),
    ("sharing:\n  timeSlicing:\n    resources:\n    - name: gpu\n      devices: some\n      replicas: 2\n", "the only valid string input is 'all'"),
    ("sharing:\n  timeSlicing:\n    resources:\n    - name: gpu\n      devices: 0\n      replicas: 2\n", "a count of devices must be > 0"),
    ("sharing:\n  timeSlicing:\n    resources:\n    - name: gpu\n      devices: [gpu0]\n      replicas: 2\n", "unsupported type for device in devices list"),
    ("flags:\n  failOnInitError: \"true\"\n", "cannot unmarshal string"),
    ("flags:\n  gfd:\n    sleepInterval: soon\n", "invalid duration"),
    ("resources:\n  gpus:\n  - name: x\n", "resources must have a 'pattern' field set"),
    ("flags: [1, 2\n", "yaml"),
])
def test_config_file_errors(doc, msg, tmp_path):
    cfgfile = os.path.join(str(tmp_path), "c.yaml")
    open(cfgfile, "w").write(doc)
    rc, _, se = run_cli(["--oneshot", "--config-file", cfgfile, "-o", os.path.join(str(tmp_path), "f")], env=base_env(tmp_path))
    assert rc == 1 and "Error: unable to load config: unable to finalize config: unable to parse config file: error parsing config file:" in se
    assert msg in se, se


def test_missing_config_file(tmp_path):
    rc, _, se = run_cli(["--oneshot", "--config-file", "/no/such/file"], env=base_env(tmp_path))
    assert rc == 1 and "error opening config file: open /no/such/file: No such file or directory" in se


# ---- daemon loop, signals, file life-cycle ----------------------------------------------------------------------
def spawn(tmp_path, args, node="c2_1xB200"):
    env = dict(os.environ)
    for k in list(env):
        if k.startswith("GFD_"):
            del env[k]
    env.update(base_env(tmp_path, node))
    env["LD_LIBRARY_PATH"] = B.FAKE_NVML_DIR + ":" + env.get("LD_LIBRARY_PATH", "")
    return subprocess.Popen([binary_path()] + args, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)


def wait_for(path, timeout=10.0):
    t0 = time.time()
    while time.time() - t0 < timeout:
        if os.path.exists(path) and os.path.getsize(path) > 0:
            return True
        time.sleep(0.02)
    return False


def test_sleep_loop_rewrites_file_with_same_timestamp_and_sigterm_removes_it(tmp_path):
    """TestRunSleep (main_test.go:182-271): file rewritten every interval, timestamp label unchanged; a non-HUP signal ends
    the loop and the deferred removeOutputFile deletes the file and gfd-tmp (main.go:149-156,212-232)."""
    out = os.path.join(str(tmp_path), "features.d", "gfd")
    os.makedirs(os.path.dirname(out))
    p = spawn(tmp_path, ["--sleep-interval", "300ms", "-o", out])
    try:
        assert wait_for(out)
        first = open(out).read()
        ino1, m1 = os.stat(out).st_ino, os.stat(out).st_mtime_ns
        deadline = time.time() + 10
        while time.time() < deadline and os.stat(out).st_ino == ino1 and os.stat(out).st_mtime_ns == m1:
            time.sleep(0.05)
        assert (os.stat(out).st_ino, os.stat(out).st_mtime_ns) != (ino1, m1), "file was not rewritten"
        second = open(out).read()
        assert first == second and re.search(r"nvidia\.com/gfd\.timestamp=[0-9]{10}\n", first)
        p.send_signal(signal.SIGTERM)
        so, se = p.communicate(timeout=10)
    finally:
        if p.poll() is None:
            p.kill()
    assert p.returncode == 0
    assert 'Received signal "terminated", shutting down.' in se and "Sleeping for 300ms" in se and se.rstrip().endswith("Exiting")
    assert not os.path.exists(out) and not os.path.exists(os.path.join(os.path.dirname(out), "gfd-tmp"))


def test_sighup_reloads_config_and_rebuilds_manager(tmp_path):
    """main.go:117-145,201-203: SIGHUP -> run() returns restart=true -> file removed -> config reloaded, new manager."""
    out = os.path.join(str(tmp_path), "gfd")
    p = spawn(tmp_path, ["--sleep-interval", "1h", "--no-timestamp", "-o", out])
    try:
        assert wait_for(out)
        # swap the node under the daemon, then HUP
        use_node(NODES["c3_8xB200"], tmp_path)
        p.send_signal(signal.SIGHUP)
        deadline = time.time() + 10
        while time.time() < deadline and not (os.path.exists(out) and "gpu.count=8" in open(out).read()):
            time.sleep(0.05)
        text = open(out).read()
        p.send_signal(signal.SIGINT)
        so, se = p.communicate(timeout=10)
    finally:
        if p.poll() is None:
            p.kill()
    assert "nvidia.com/gpu.count=8\n" in text
    assert se.count("Loading configuration.") == 2 and "Received SIGHUP, restarting." in se and 'Received signal "interrupt"' in se
    assert p.returncode == 0 and not os.path.exists(out)


def test_oneshot_keeps_file_and_leaves_tmp_dir(tmp_path):
    out = os.path.join(str(tmp_path), "gfd")
    rc, _, se = run_cli(["--oneshot", "-o", out], env=base_env(tmp_path))
    assert rc == 0 and os.path.exists(out) and os.path.isdir(os.path.join(str(tmp_path), "gfd-tmp"))


def test_unwritable_output_directory_is_an_error(tmp_path):
    rc, _, se = run_cli(["--oneshot", "-o", "/proc/nope/gfd"], env=base_env(tmp_path))
    assert rc == 1 and "error writing file '/proc/nope/gfd'" in se


def test_daemon_memory_does_not_grow_with_the_number_of_passes(tmp_path):
    """A DaemonSet pod runs for months: every Init..Shutdown bracket must give back what it took (device objects, cached
    answers, the dlopen handle).  Peak RSS after 1500 passes of a 56-MIG-device node must equal the peak after 30, give or
    take allocator noise.  (`make asan` + detect_leaks over 200 passes is the byte-exact version of this check.)"""
    if "san" in os.environ.get("GFD_TEST_BINDIR", ""):
        pytest.skip("sanitizer builds keep freed memory in quarantine: RSS says nothing there (LSan checks leaks instead)")
    fix = use_node(N.nodes()["c4_single_8x7x1g23"], str(tmp_path))
    empty = os.path.join(str(tmp_path), "pci")
    os.makedirs(empty)
    peaks = []
    for passes in (30, 1500):
        script = ("import resource, subprocess, sys; "
                  "p = subprocess.run(sys.argv[1:], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL); "
                  "print(p.returncode, resource.getrusage(resource.RUSAGE_CHILDREN).ru_maxrss)")
        env = dict(os.environ, LD_LIBRARY_PATH=B.FAKE_NVML_DIR + ":" + os.environ.get("LD_LIBRARY_PATH", ""), GFD_FAKE_NVML=fix,
                   GFD_B200_PCI_DEVICES_ROOT=empty)
        out = subprocess.run(["python3", "-c", script, binary_path(), "--passes", str(passes), "--sleep-interval", "0s", "--mig-strategy=single",
                              "--no-timestamp", "-o", os.path.join(str(tmp_path), "out")], env=env, capture_output=True, text=True, timeout=300)
        rc, kb = out.stdout.split()
        assert rc == "0", out.stderr
        peaks.append(int(kb))
    assert peaks[1] - peaks[0] < 1024, "peak RSS grew by %d KiB between 30 and 1500 passes" % (peaks[1] - peaks[0])


MANIFESTS = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "manifests.json")))["manifests"]


@pytest.mark.parametrize("rel", sorted(MANIFESTS))
def test_started_the_way_the_reference_manifests_start_it(rel, tmp_path):
    """Every manifest the reference ships (static DaemonSets, the Job template) starts the binary with some args + env:
    the drop-in must accept exactly those, come up with the strategy they ask for, and write where they mount features.d."""
    m = MANIFESTS[rel]
    fix = use_node(N.nodes()["c4_single_8x7x1g23"], str(tmp_path))
    out = os.path.join(str(tmp_path), "gfd")
    env = dict(m["env"], GFD_FAKE_NVML=fix, GFD_OUTPUT_FILE=out, GFD_NO_TIMESTAMP="true")
    args = list(m["args"])
    if "--oneshot" not in args:
        env["GFD_ONESHOT"] = "true"  # the DaemonSets loop for ever; one pass is enough here
    rc, _, se = run_cli(args, env=env)
    assert rc == 0, se
    cfg = json.loads(re.search(r"Running with config:\n(\{.*?\n\})\n", se, re.S).group(1))
    assert cfg["flags"]["migStrategy"] == m["env"].get("MIG_STRATEGY", "none")
    labels = dict(l.split("=", 1) for l in open(out).read().splitlines())
    assert labels.get("nvidia.com/mig.strategy", "none") == m["env"].get("MIG_STRATEGY", "none") or m["env"].get("MIG_STRATEGY", "none") == "none"
    # the default output file lives in the directory the manifests mount for NFD
    rc, so, _ = run_cli(["--help"])
    default_out = re.search(r"--output-file.*?\(default: \"?([^\")]+)\"?\)", so, re.S).group(1)
    assert os.path.dirname(default_out) in [p.rstrip("/") for p in m["mount_paths"]]
