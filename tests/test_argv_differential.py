"""Command line + environment of the binary vs oracle/config_oracle.py on GENERATED invocations (hypothesis).

Each example picks a value for some of the reference's flags and delivers every one of them in a randomly chosen way
-- on the command line (one or two dashes, `=value` or a separate argument, any alias, any order, bools as bare flag
or `=true/=0/=F`...), through one of the flag's environment variables, or through both (the command line must win;
with two variables set the FIRST one listed in main.go:33-82 must win) -- plus, sometimes, something that must be
rejected (unknown flag, two spellings of one flag, a value that does not parse, a missing argument).  The config the
binary logs must be the one the oracle derives from the chosen values (cli.Context.IsSet semantics), or the run must
end with exit 1 and no config."""
import json
import os
import re

from hypothesis import HealthCheck, given, settings, strategies as st

import nodes as N
from oracle import config_oracle as C
from shimutil import run_cli, use_node

# flag -> (oracle key, kind, spellings, environment variables in the reference's order)
FLAGS = {
    "mig-strategy": ("migStrategy", "str", ["mig-strategy"], ["GFD_MIG_STRATEGY", "MIG_STRATEGY"]),
    "fail-on-init-error": ("failOnInitError", "bool", ["fail-on-init-error"], ["GFD_FAIL_ON_INIT_ERROR", "FAIL_ON_INIT_ERROR"]),
    "no-timestamp": ("noTimestamp", "bool", ["no-timestamp"], ["GFD_NO_TIMESTAMP"]),
    "sleep-interval": ("sleepInterval", "dur", ["sleep-interval"], ["GFD_SLEEP_INTERVAL"]),
    "machine-type-file": ("machineTypeFile", "str", ["machine-type-file"], ["GFD_MACHINE_TYPE_FILE"]),
}
TRUE, FALSE = ["1", "t", "T", "TRUE", "true", "True"], ["0", "f", "F", "FALSE", "false", "False"]
DURS = ["1s", "90s", "1h2m3s", "1.5s", "250ms", "2m", "100ns", "1h", "0s", "10us", "+5s"]
STRS = ["none", "single", "mixed", "bogus", "", "/dev/null", "a b", "x=y", "-dash", "--two"]


@st.composite
def invocation(draw):
    argv, env, want, bad = [], {}, {}, False
    for flag in draw(st.lists(st.sampled_from(sorted(FLAGS)), unique=True, max_size=5)):
        key, kind, spellings, envs = FLAGS[flag]
        if kind == "bool":
            val = draw(st.booleans())
            text = draw(st.sampled_from(TRUE if val else FALSE))
        elif kind == "dur":
            text = draw(st.sampled_from(DURS))
            val = C.parse_duration(text)
        else:
            text = val = draw(st.sampled_from(STRS))
        how = draw(st.sampled_from(["cli", "cli", "env", "both", "env2"]))
        if how in ("env", "both", "env2"):
            if how == "env2" and len(envs) > 1:      # both variables set: the first listed wins
                env[envs[0]] = text
                env[envs[1]] = draw(st.sampled_from(FALSE + TRUE)) if kind == "bool" else ("7s" if kind == "dur" else "loser")
            else:
                env[draw(st.sampled_from(envs))] = text
            want[key] = val
        if how in ("cli", "both"):
            if how == "both":  # the command line wins over the environment: give it a DIFFERENT value
                if kind == "bool":
                    val = not val
                    text = draw(st.sampled_from(TRUE if val else FALSE))
                elif kind == "dur":
                    text = draw(st.sampled_from([d for d in DURS if C.parse_duration(d) != val]))
                    val = C.parse_duration(text)
                else:
                    text = val = val + "-cli"
            dashes = draw(st.sampled_from(["-", "--"]))
            name = dashes + draw(st.sampled_from(spellings))
            if kind == "bool":
                form = draw(st.sampled_from(["bare", "eq"])) if val else "eq"
                argv.append([name] if form == "bare" else [name + "=" + text])
            elif draw(st.booleans()):
                argv.append([name + "=" + text])
            else:
                argv.append([name, text])
            want[key] = val
    # the output file: any of its three spellings, sometimes twice
    out_spellings = draw(st.lists(st.sampled_from(["output-file", "output", "o"]), min_size=1, max_size=2))
    for sp in out_spellings:
        argv.append([draw(st.sampled_from(["-", "--"])) + sp, "@OUT@"])
    if len(set(out_spellings)) > 1:
        bad = True  # normalizeFlags (V/cli/flag.go:164-195): "Cannot use two forms of the same flag"
    garbage = draw(st.sampled_from([None, None, None, ["--no-such-flag"], ["--sleep-interval=soon"], ["--no-timestamp=maybe"], ["---x"], ["--=x"], ["-sleep-interval"]]))
    order = draw(st.permutations(argv))
    flat = [a for group in order for a in group]
    if garbage is not None:
        pos = draw(st.integers(0, len(order)))
        flat = [a for group in order[:pos] for a in group] + garbage + [a for group in order[pos:] for a in group]
        bad = True
        if garbage == ["-sleep-interval"] and pos < len(order):
            # the next argument is swallowed as the (unparsable or parsable) duration: outcome depends on it -- keep it simple
            flat = [a for group in order for a in group] + garbage   # at the very end: "flag needs an argument"
    return flat, env, want, bad


@settings(max_examples=400, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow])
@given(inv=invocation())
def test_argv_and_env_match_oracle(inv, tmp_path_factory):
    flat, env, want, bad = inv
    tmp = str(tmp_path_factory.mktemp("argv"))
    out = os.path.join(tmp, "gfd")
    fix = use_node(N.nodes()["zero_gpus"], tmp)
    empty = os.path.join(tmp, "pci")
    os.makedirs(empty)
    args = ["--oneshot"] + [out if a == "@OUT@" else a for a in flat]
    e = dict(env, GFD_FAKE_NVML=fix, GFD_B200_PCI_DEVICES_ROOT=empty)
    rc, _, se = run_cli(args, env=e)
    m = re.search(r"Running with config:\n(\{.*?\n\})\n", se, re.S)
    if bad:
        assert rc == 1 and m is None, (args, env, se[-500:])
        return
    cli_set = dict(want, oneshot=True, outputFile=out)
    expect = C.load_config(False, None, cli_set)
    assert m is not None, (args, env, se[-500:])
    assert json.loads(m.group(1)) == expect, (args, env)
    assert rc == 0, (args, env, se[-500:])
