"""Run bench.py with the given env overrides and print a one-line digest (tuning helper)."""
import json, os, subprocess, sys
env = dict(os.environ)
args = []
for a in sys.argv[1:]:
    if "=" in a and not a.startswith("-"):
        k, v = a.split("=", 1); env[k] = v
    else:
        args.append(a)
out = subprocess.run([sys.executable, "bench.py", "--no-cpu-baseline", "--no-e2e"] + args, env=env,
                     capture_output=True, text=True)
try:
    j = json.loads(out.stdout.strip().splitlines()[-1])
    k = j["kernels"]
    print(" ".join(sys.argv[1:]), "| ms/step %.2f tok/s %.0f frac_burst %.3f clk %s %s |" % (
        j["ms_per_step"], j["value"], j["step_frac_of_bf16_peak"]["burst"], j["clocks"]["sm_mhz"], j["clocks"]["reasons"]),
        " ".join("%s %.2f" % (n, k[n]["ms_per_step"]) for n in ("lse_fwd", "dlogits", "softmax_grad", "dW", "dX", "scatter", "aux") if n in k))
except Exception as e:
    print("FAILED", sys.argv[1:], e, out.stdout[-500:], out.stderr[-1500:])
