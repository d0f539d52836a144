"""Driver around the solvers: the reference's `code/main.py` for the five primal first-order solvers,
with the iterate, the ground truth and every metric kept on the device.

Reference behaviour mirrored here (code/main.py):
  :37-51    gen_data(seed)             -> glprox.dataset.gen_data (fixture) / gen_data_device (at scale)
  :113-130  solve_routine              -> one log line per solver, "[name      ]: cpu: .., iter: .., ..."
  :94-110   write_to_table             -> the Markdown "Statistics" table of doc/report.md:439-447
  :48-50    errfun / errfun_exact / sparsity -> GroupLassoProblem.metrics (device reductions)
The CVX / Mosek / Gurobi rows and the SVG plots of the reference are outside the hot path (no such
solvers or plotting stack here): the "err-to-cvx-*" columns are replaced by one "err-to-best" column,
the distance errfun(x_best, x) to the run with the lowest objective.

  python main.py                               # the homework fixture (seed 97006855, fp64), all five solvers
  python main.py --m 8192 --n 32768 --dtype f32 --solvers FProxGD,ProxGD
  python main.py --m 8192 --n 32768 --save /tmp/p      # also store the problem as GLPSHARD files
  python main.py --load /tmp/p                 # solve a stored problem
  torchrun --nproc-per-node 8 main.py --rows 32768 --cols 131072 --dtype f32   # one column shard per GPU
"""
import argparse
import logging
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

SOLVERS = {  # name -> (row label of doc/report.md:439-447, algo key)
    "SGD": ("SGD Primal", "SGD"),
    "GD": ("GD Primal", "GD"),
    "FGD": ("FGD Primal", "FGD"),
    "ProxGD": ("ProxGD Primal", "ProxGD"),
    "FProxGD": ("FProxGD Primal", "FProxGD"),
    "ALM_dual": ("ALM Dual", "ALM_dual"),         # doc/report.md:570
    "ADMM_dual": ("ADMM Dual", "ADMM_dual"),      # doc/report.md:571
    "ADMM_primal": ("ADMM Primal", "ADMM_primal"),   # doc/report.md:572
}
HEADERS = ["solver", "cpu", "iter", "optval", "sparsity", "err-to-exact", "err-to-best"]


def format_row(name, seconds, iters, fval, met, err_best):
    """The value formats of code/main.py:117-123."""
    return {"solver": name, "cpu": "%5.2f" % seconds, "iter": "%5d" % (-1 if iters is None else iters),
            "optval": "%6.5E" % fval, "sparsity": "%6.4f" % met["sparsity"],
            "err-to-exact": "%3.2E" % met["err_exact"],
            "err-to-best": "n/a" if err_best is None else "%3.2E" % err_best}


def log_line(row):
    """code/main.py:126-128."""
    items = [(k, v) for k, v in row.items() if k != "solver"]
    return ("[%-10s]: " % row["solver"]) + ", ".join(k + ": " + v for k, v in items)


def markdown_table(rows, title="Statistics"):
    """A GitHub-flavoured table laid out like pytablewriter's MarkdownTableWriter (code/main.py:94-110)."""
    cols = [[h] + [r[h] for r in rows] for h in HEADERS]
    widths = [max(3, max(len(c) for c in col)) for col in cols]
    line = lambda cells: "|" + "|".join(" " + c.ljust(w) + " " for c, w in zip(cells, widths)) + "|"
    out = ["# " + title, line(HEADERS), "|" + "|".join("-" * (w + 2) for w in widths) + "|"]
    out += [line([r[h] for h in HEADERS]) for r in rows]
    return "\n".join(out)


def parse_args(argv=None):
    p = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    p.add_argument("--seed", type=int, default=97006855)
    # (--rows / --cols are the spellings to use under torchrun, whose own parser claims the prefix "--m")
    p.add_argument("--m", "--rows", dest="m", type=int, default=0, help="rows of A (0 = the reference fixture 256 x 512)")
    p.add_argument("--n", "--cols", dest="n", type=int, default=0, help="columns of A = row groups of x")
    p.add_argument("--l", "--rhs", dest="l", type=int, default=2, help="right-hand sides = columns of x")
    p.add_argument("--mu", type=float, default=1e-2)
    p.add_argument("--dtype", choices=["f32", "f64"], default="f64")
    p.add_argument("--solvers", default="SGD,GD,FGD,ProxGD,FProxGD")
    p.add_argument("--maxit", type=int, default=0, help="override opts['maxit'] (per continuation stage)")
    p.add_argument("--alpha0", type=float, default=0.0, help="override opts['alpha0'] (0: reference default; "
                   "-1: 1/(sqrt(m)+sqrt(n))^2, a safe step for N(0,1) matrices of any size)")
    p.add_argument("--save", default="", help="store the generated problem under this path prefix")
    p.add_argument("--load", default="", help="solve the problem stored under this path prefix")
    p.add_argument("--log-level", default="INFO")
    return p.parse_args(argv)


def build_problem(args, rank, world):
    """-> (A, b, x0, u, mu, n_total) as CUDA tensors holding this rank's shard."""
    import torch
    from glprox import dataset
    if args.load:
        A, b, x0, u, mu, _ = dataset.load_problem_device(args.load, rank, world)
        return A, b, x0, u, mu, dataset.open_shards(args.load)[0].n_total
    if args.m and args.n:
        A, b, x0, u, mu, _ = dataset.gen_data_device(args.m, args.n, args.l, args.seed, args.dtype, rank, world, args.mu)
        return A, b, x0, u, mu, args.n
    n, m, l, mu, A, b, u, x0 = dataset.gen_data(args.seed)
    tdt = torch.float32 if args.dtype == "f32" else torch.float64
    if world > 1:
        from glprox import shard_bounds
        c0, c1 = shard_bounds(n, world, 128 if args.dtype == "f32" else 64)[rank]
        A, u, x0 = A[:, c0:c1], u[c0:c1], x0[c0:c1]
    dev = lambda z: torch.from_numpy(np.ascontiguousarray(z)).to("cuda", tdt)
    return dev(A), dev(b), dev(x0), dev(u), mu, n


def main(argv=None):
    args = parse_args(argv)
    logging.basicConfig(level=getattr(logging, args.log_level.upper()), format="%(message)s")
    logger = logging.getLogger("opt")
    import torch
    from glprox import GroupLassoProblem, attach_distributed, dataset
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
        dist.init_process_group("nccl")
    A, b, x0, u, mu, n_total = build_problem(args, rank, world)
    if args.save:
        c0 = 0
        if world > 1:
            from glprox import shard_bounds
            c0 = shard_bounds(n_total, world, 128 if A.dtype == torch.float32 else 64)[rank][0]
        path = dataset.write_shard(args.save, rank, world, A, b, x0, u, n_total=n_total, col0=c0, mu=mu, seed=args.seed)
        logger.info("stored %s", path)
    prob = GroupLassoProblem(A, b)
    if world > 1:
        attach_distributed(prob, n_total)
    opts = {}
    if args.maxit:
        opts["maxit"] = args.maxit
    if args.alpha0:
        m = A.shape[0]
        opts["alpha0"] = args.alpha0 if args.alpha0 > 0 else 1.0 / (np.sqrt(m) + np.sqrt(n_total)) ** 2
    runs = []
    for key in [s.strip() for s in args.solvers.split(",") if s.strip()]:
        if key not in SOLVERS:
            raise SystemExit("unknown solver %r (choose from %s)" % (key, ", ".join(SOLVERS)))
        label, algo = SOLVERS[key]
        if algo in ("ADMM_dual", "ADMM_primal", "ALM_dual"):
            fn = {"ADMM_dual": prob.solve_admm_dual, "ADMM_primal": prob.solve_admm_primal, "ALM_dual": prob.solve_alm_dual}[algo]
            x, iters, out = fn(x0, mu, {"maxit": args.maxit} if args.maxit else {})
        else:
            x, iters, out = prob.solve(algo, x0, mu, opts)      # x stays in HBM
        met = prob.metrics(x, u, mu)
        runs.append((label, out["tt"], iters, float(out["fval"]), met, x))
    best = min(runs, key=lambda r: r[3])[5] if runs else None
    rows = []
    for label, secs, iters, fval, met, x in runs:
        err_best = prob.metrics(x, best, mu)["err_ref"] if len(runs) > 1 else None
        rows.append(format_row(label, secs, iters, fval, met, err_best))
        if rank == 0:
            logger.info(log_line(rows[-1]))
    if rank == 0:
        print(markdown_table(rows))
    prob.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return rows


if __name__ == "__main__":
    main()
