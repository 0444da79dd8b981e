"""Command-line twin of the reference's SSPS/ssps_wrapper.jl (argument table: SSPS/src/SSPS.jl:48-106).

    python -m ssps_b200.ssps_wrapper <timeseries> <ref_graph> <output.json> <timeout> [flags]

Same positionals, same flags, same per-chain JSON (update_results_storediff + SSPS.jl:223-230), so
run_ssps/Snakefile, postprocess_samples.jl and scoring.jl consume the output unchanged.  The
sampling itself runs in libssps_b200.so on one B200; there is no CPU path.

Extensions (not in the reference): --n-chains K writes K chains from one process
(<output> gets the first chain, or use a `{chain}` placeholder in the output path); --seed;
--device; --gpus N shards the K chains over N GPUs of the box (one process, ssps_run_multi; the
pooled statistics are all-reduced by the library's own NCCL communicator); --posterior-json writes
the pooled postprocess_samples.jl-style summary directly; --no-chain-json skips the per-chain files
(thousands of chains).
"""
from __future__ import annotations

import argparse
import json
import math
import sys
import time

import numpy as np

from .preprocess import load_formatted_data
from .sampler import Sampler

_INT64_MAX = 2 ** 63 - 1


def parse_script_arguments(argv=None) -> argparse.Namespace:
    p = argparse.ArgumentParser(prog="ssps_wrapper", description="SSPS on B200 (drop-in for ssps_wrapper.jl)")
    p.add_argument("timeseries_filename", help="name of timeseries data file")
    p.add_argument("ref_graph_filename", help="name of reference graph data file")
    p.add_argument("output_path", help="name of output JSON file")
    p.add_argument("timeout", type=float, nargs="?", default=3600.0,
                   help="execution timeout (in seconds). When reached, terminate and output results as they are.")
    p.add_argument("--thinning", type=int, default=1, help="number of proposals to make for each sample taken")
    p.add_argument("--large-indeg", type=float, default=20.0,
                   help="an approximate upper bound on the number of parents for any node")
    p.add_argument("--lambda-max", type=float, default=15.0, help="hyperparameter for lambda prior")
    p.add_argument("--regression-deg", type=int, default=-1,
                   help="regression polynomial degree in Gaussian DBN (default: full)")
    p.add_argument("--lambda-prop-std", type=float, default=0.25,
                   help="standard deviation of lambda's proposal distribution")
    p.add_argument("--n-steps", type=int, default=-1, help="Terminate a markov chain after it runs this many steps.")
    p.add_argument("--proposal", default="sparse", choices=("sparse", "uniform"),
                   help="(DEPRECATED in the reference.) only 'sparse' is implemented")
    p.add_argument("--resume-sampling", default="", help="Continue sampling the markov chain stored in the given JSON file")
    p.add_argument("--multithread", action="store_true",
                   help="accepted for compatibility: every vertex already runs in parallel on the GPU")
    # extensions
    p.add_argument("--n-chains", type=int, default=1)
    p.add_argument("--seed", type=int, default=None)
    p.add_argument("--device", type=int, default=-1)
    p.add_argument("--gpus", type=int, default=1, help="shard --n-chains over this many GPUs of the box (devices 0..N-1)")
    p.add_argument("--no-chain-json", action="store_true", help="do not write the per-chain change-point JSON files")
    p.add_argument("--posterior-json", default="", help="also write the pooled edge posterior / PSRF summary here")
    p.add_argument("--burnin", type=float, default=0.5)
    return p.parse_args(argv)


def recover_state(samples: dict):
    """recover_state! (SSPS.jl:24-43): last lambda of every vertex; parents whose last change-point value is 1."""
    lam = [svec[-1][1] for svec in samples["lambda"]]
    psets = []
    for pdict in samples["parent_sets"]:
        psets.append(sorted(int(k) - 1 for k, v in pdict.items() if v[-1][1] == 1))
    return psets, lam


def perform_inference(timeseries_filename, ref_graph_filename, output_path, timeout, n_steps, thinning,
                      large_indeg, lambda_max, regression_deg, lambda_prop_std, proposal, resume_sampling,
                      multithread, n_chains=1, seed=None, device=-1, posterior_json="", burnin=0.5, gpus=1,
                      no_chain_json=False):
    """perform_inference (SSPS.jl:135-233)."""
    if proposal not in ("sparse", "uniform"):
        raise SystemExit("--proposal must be 'sparse' or 'uniform' (SSPS.jl:87-88)")
    if multithread:
        proposal = "sparse"             # "This flag overrides the --proposal argument, setting it to 'sparse'" (SSPS.jl:98)
    if proposal == "uniform":
        if regression_deg >= 2:
            raise SystemExit("--proposal uniform (deprecated in the reference, SSPS.jl:88) is implemented for --regression-deg <= 1 only")
        if gpus > 1:
            raise SystemExit("--proposal uniform runs on one GPU (one thread per chain)")
    if regression_deg < 1:
        print("note: --regression-deg %d builds an EMPTY design in the reference (construct_B, dbn_distributions.jl:179-194): "
              "the likelihood is flat and only the edge prior acts.  Reproduced as is; every shipped config passes 1."
              % regression_deg, file=sys.stderr)
    series, prior, _ = load_formatted_data(timeseries_filename, ref_graph_filename)
    print("Invoking SSPS on input files:\n\t%s\n\t%s" % (timeseries_filename, ref_graph_filename))
    t_start = time.time()
    kw = dict(regression_deg=regression_deg, large_indeg=large_indeg, lambda_max=lambda_max,
              lambda_prop_std=lambda_prop_std, record_changepoints=not no_chain_json, proposal=proposal)
    if seed is None:
        seed = int(time.time_ns() & 0x7FFFFFFFFFFFFFFF)   # the reference never seeds its RNG either
    multi = gpus > 1
    if multi:
        from .multi_gpu import MultiGpuSampler
        if n_chains < gpus:
            raise SystemExit("--gpus %d needs at least that many chains (--n-chains %d)" % (gpus, n_chains))
        smp = MultiGpuSampler(n_chains, list(range(gpus)), seed=seed, **kw)
    else:
        smp = Sampler(device=device, **kw)
    smp.set_timeseries(series, prior)
    init_sets = init_lam = None
    if resume_sampling:
        print("Resuming Markov Chain stored at", resume_sampling)
        with open(resume_sampling) as f:
            psets, lam = recover_state(json.load(f))
        init_sets, init_lam = [psets] * n_chains, [lam] * n_chains
    if multi:
        smp.generate(init_parent_sets=init_sets, init_lambda=init_lam)
    elif init_sets is not None:
        smp.generate(n_chains, seed=seed, init_parent_sets=init_sets, init_lambda=init_lam)
    else:
        smp.generate(n_chains, seed=seed)

    if n_steps > 0:
        print("Sampling for %s seconds. (Or %d iterations.)" % (timeout, n_steps))
        # `while prop_count <= n_steps` with `thinning` sweeps per pass (mcmc_inference.jl:58-61, quirk Q3)
        total = (n_steps // thinning + 1) * thinning
    else:
        print("Sampling for %s seconds. (No iteration limit.)" % timeout)
        total = _INT64_MAX // (2 * thinning) * thinning
    done, batch = 0, 1000 * thinning
    while done < total:
        left = timeout - (time.time() - t_start)
        if left <= 0.0:
            break
        step = min(batch, total - done)
        got = smp.update_loop(step, thinning=thinning, timeout=left)
        done += got
        if got < step:
            break
    t_elapsed = time.time() - t_start

    print("Saving results to JSON file:")
    for c in range(n_chains if not no_chain_json else 0):
        path = output_path.replace("{chain}", str(c)) if "{chain}" in output_path else (
            output_path if c == 0 else "%s.chain%d" % (output_path, c))
        with open(path, "w") as f:
            json.dump(smp.results(c, t_elapsed=t_elapsed), f)
        print("\t" + path)
    if posterior_json:
        from .postprocess import summarize_sampler
        with open(posterior_json, "w") as f:
            json.dump(summarize_sampler(smp, burnin=burnin, t_elapsed=t_elapsed), f)
        print("\t" + posterior_json)
    smp.close()
    return 0


def main(argv=None) -> int:
    a = parse_script_arguments(argv)
    return perform_inference(a.timeseries_filename, a.ref_graph_filename, a.output_path, a.timeout, a.n_steps,
                             a.thinning, a.large_indeg, a.lambda_max, a.regression_deg, a.lambda_prop_std,
                             a.proposal, a.resume_sampling, a.multithread, n_chains=a.n_chains, seed=a.seed,
                             device=a.device, posterior_json=a.posterior_json, burnin=a.burnin, gpus=a.gpus,
                             no_chain_json=a.no_chain_json)


if __name__ == "__main__":
    sys.exit(main())
