"""Worker functions for the multi-process tests (must live in an importable module for mp 'spawn')."""
import copy
import os
import sys
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def parity_worker(rank, world_size, schedule_type, n_layers, v, m, results_queue, master_port=None, optimizer=None,
                  dp=1, device="cpu", rule="reference", costs=None):
    """Pipelined loss + per-parameter grads must equal the single-process full-model ones (same seed, dropout 0)."""
    try:
        import torch
        import torch.distributed as dist
        from ppb200.models import ModelArgs, Transformer
        from ppb200.api import manual_model_split, make_schedule
        from ppb200.parallel.mesh import init_mesh
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(master_port), RANK=str(rank), WORLD_SIZE=str(world_size))
        torch.set_num_threads(1)
        dist.init_process_group("gloo", rank=rank, world_size=world_size)
        mesh = init_mesh(dp_size=dp, pp_size=world_size // dp)
        P = mesh.pp_size
        args = ModelArgs(dim=64, n_layers=n_layers, n_heads=4, vocab_size=101, ffn_dim=96, dropout=0.0)
        if costs is not None:   # force the positions of the sub-layer cuts through "measured" unit costs
            from ppb200.parallel.partition import set_measured_costs
            set_measured_costs(args, 1024, costs)
        torch.manual_seed(1234)
        full = Transformer(args)
        B, S = 8 * dp, 16
        g = torch.Generator().manual_seed(99)
        x = torch.randint(0, 101, (B, S), generator=g)
        y = torch.randint(0, 101, (B, S), generator=g)
        # ---- single-process oracle: mean over micro-batches of the mean token CE (== plain mean for equal chunks)
        oracle = copy.deepcopy(full)
        logits = oracle(x)
        loss_ref = torch.nn.functional.cross_entropy(logits.reshape(-1, 101), y.reshape(-1))
        loss_ref.backward()
        ref_grads = {n: p.grad.clone() for n, p in oracle.named_parameters()}
        # ---- pipelined
        num_stages = P * v
        stages = []
        for i in range(v):
            mdl = copy.deepcopy(full)
            st = manual_model_split(mdl, mesh.pp_idx + P * i, num_stages, torch.device("cpu"), rule=rule)
            st.dp_group = mesh.dp_group
            stages.append(st)
        sched = make_schedule(schedule_type, stages, m, None, pp_group=mesh.pp_group, dp_group=mesh.dp_group,
                              optimizer=None)
        Bl = B // dp
        xl, yl = x[mesh.dp_idx * Bl:(mesh.dp_idx + 1) * Bl], y[mesh.dp_idx * Bl:(mesh.dp_idx + 1) * Bl]
        losses = []
        first, last = 0 in sched.local, (num_stages - 1) in sched.local
        sched.step(*((xl,) if first else ()), target=yl if last else None, losses=losses if last else None)
        max_err = 0.0
        for st in stages:
            if dp > 1 and st.n_total:
                st.grad.div_(dp)  # the schedule's ALLREDUCE action already summed over the replicas
            for n, gview in st.grad_views().items():
                err = (gview - ref_grads[n]).abs().max().item()
                scale = ref_grads[n].abs().max().item() + 1e-8
                max_err = max(max_err, err / scale)
        out = {"rank": rank, "max_rel_grad_err": max_err}
        if last:
            lp = torch.stack([l.reshape(()) for l in losses]).mean()
            if dp > 1:
                dist.all_reduce(lp, group=mesh.dp_group)
                lp = lp / dp
            out["loss_err"] = abs(lp.item() - loss_ref.item())
            out["n_losses"] = len(losses)
        elif dp > 1:
            pass
        gathered = [None] * world_size
        dist.all_gather_object(gathered, out)
        if rank == 0:
            results_queue.put({"ranks": gathered, "bubble": sched.measured()})
        dist.barrier()
        dist.destroy_process_group()
    except Exception as e:  # noqa: BLE001
        traceback.print_exc()
        results_queue.put({"error": f"rank {rank}: {e}"})


def train_worker(rank, world_size, schedule_type, results_queue, master_port=None, steps=6):
    """A few optimizer steps must reduce the loss on a fixed batch (dropout on, AdamW on)."""
    try:
        import torch
        import torch.distributed as dist
        from ppb200.models import ModelArgs, Transformer
        from ppb200.api import manual_model_split, make_schedule
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(master_port), RANK=str(rank), WORLD_SIZE=str(world_size))
        torch.set_num_threads(1)
        dist.init_process_group("gloo", rank=rank, world_size=world_size)
        args = ModelArgs(dim=64, n_layers=4, n_heads=4, vocab_size=101, ffn_dim=96, dropout=0.1)
        v = 2 if schedule_type == "Interleaved1F1B" else 1
        stages = []
        for i in range(v):
            torch.manual_seed(7)
            st = manual_model_split(Transformer(args), rank + world_size * i, world_size * v, torch.device("cpu"))
            stages.append(st)
        sched = make_schedule(schedule_type, stages, 4, None, optimizer={"lr": 3e-3, "weight_decay": 0.0})
        g = torch.Generator().manual_seed(5)
        x = torch.randint(0, 101, (8, 16), generator=g)
        y = torch.randint(0, 101, (8, 16), generator=g)
        first, last = 0 in sched.local, (world_size * v - 1) in sched.local
        hist = []
        for _ in range(steps):
            losses = []
            sched.step(*((x,) if first else ()), target=y if last else None, losses=losses if last else None)
            if last:
                hist.append(torch.stack([l.reshape(()) for l in losses]).mean().item())
        if last:
            results_queue.put({"losses": hist})
        dist.barrier()
        dist.destroy_process_group()
    except Exception as e:  # noqa: BLE001
        traceback.print_exc()
        results_queue.put({"error": f"rank {rank}: {e}"})


def tp_parity_worker(rank, world_size, tp, n_layers, m, results_queue, master_port=None, act="gelu", norm="rmsnorm",
                     split_bw=False, rope=False, dp=1, recompute=False, postln=False):
    """(pp x tp) mesh, pre-LN block: loss and per-parameter gradients (each rank's SHARD of them) must equal the
    single-process full-model ones."""
    try:
        import torch
        import torch.distributed as dist
        from ppb200.models import ModelArgs, Transformer
        from ppb200.models.tp import shard_model_for_tp, shard_tensor
        from ppb200.api import make_schedule
        from ppb200.parallel.partition import prune_to_stage
        from ppb200.parallel.stage import PipelineStage
        from ppb200.parallel.mesh import init_mesh
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(master_port), RANK=str(rank), WORLD_SIZE=str(world_size))
        torch.set_num_threads(1)
        dist.init_process_group("gloo", rank=rank, world_size=world_size)
        mesh = init_mesh(dp_size=dp, pp_size=world_size // (tp * dp), tp_size=tp)
        P = mesh.pp_size
        args = ModelArgs(dim=64, n_layers=n_layers, n_heads=4, vocab_size=101, ffn_dim=96, dropout=0.0, causal=True,
                         cross_attn=False, activation=act, norm_first=True, norm=norm, rope=rope)
        if postln:   # the reference's own block: post-LN, self + cross attention
            args = ModelArgs(dim=64, n_layers=n_layers, n_heads=4, vocab_size=101, ffn_dim=96, dropout=0.0, activation=act,
                             norm=norm)
        torch.manual_seed(4321)
        full = Transformer(args)
        B, S = 8 * dp, 16
        g = torch.Generator().manual_seed(7)
        x = torch.randint(0, 101, (B, S), generator=g)
        y = torch.randint(0, 101, (B, S), generator=g)
        oracle = copy.deepcopy(full)
        loss_ref = torch.nn.functional.cross_entropy(oracle(x).reshape(-1, 101), y.reshape(-1))
        loss_ref.backward()
        ref_grads = {n: p.grad.clone() for n, p in oracle.named_parameters()}
        mdl = copy.deepcopy(full)
        prune_to_stage(mdl, mesh.pp_idx, P)
        shard_model_for_tp(mdl, mesh.tp_idx, tp)
        st = PipelineStage(mdl, mesh.pp_idx, P, torch.device("cpu"), tp_group=mesh.tp_group, dp_group=mesh.dp_group)
        custom = None
        if split_bw:
            from ppb200.parallel.schedule import zero_bubble_order
            custom = zero_bubble_order(P, m)
        sched = make_schedule("1F1B", st, m, None, pp_group=mesh.pp_group, pp_group_bwd=mesh.pp_group_bwd, optimizer=None,
                              actions=custom, dp_group=mesh.dp_group, recompute=recompute)
        losses = []
        first, last = 0 in sched.local, (P - 1) in sched.local
        Bl = B // dp
        xl, yl = x[mesh.dp_idx * Bl:(mesh.dp_idx + 1) * Bl], y[mesh.dp_idx * Bl:(mesh.dp_idx + 1) * Bl]
        sched.step(*((xl,) if first else ()), target=yl if last else None, losses=losses if last else None)
        if dp > 1 and st.n_total:
            st.grad.div_(dp)      # the ALLREDUCE action summed over the replicas
        max_err = 0.0
        worst = ""
        for n, gview in st.grad_views().items():
            want = shard_tensor(n, ref_grads[n], args.n_heads, mesh.tp_idx, tp, act == "swiglu")
            err = (gview - want).abs().max().item() / (want.abs().max().item() + 1e-8)
            if err > max_err:
                max_err, worst = err, n
        out = {"rank": rank, "max_rel_grad_err": max_err, "worst": worst}
        if last:
            lp = torch.stack([l.reshape(()) for l in losses]).mean()
            if dp > 1:
                dist.all_reduce(lp, group=mesh.dp_group)
                lp = lp / dp
            out["loss_err"] = abs(lp.item() - loss_ref.item())
        gathered = [None] * world_size
        dist.all_gather_object(gathered, out)
        if rank == 0:
            results_queue.put({"ranks": gathered})
        dist.barrier()
        dist.destroy_process_group()
    except Exception as e:  # noqa: BLE001
        traceback.print_exc()
        results_queue.put({"error": f"rank {rank}: {e}"})


def clip_worker(rank, world_size, results_queue, master_port=None, max_norm=0.05):
    """Global-norm gradient clipping across pipeline ranks == torch.nn.utils.clip_grad_norm_ + AdamW on the full model."""
    try:
        import torch
        import torch.distributed as dist
        from ppb200.models import ModelArgs, Transformer
        from ppb200.api import manual_model_split, make_schedule
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(master_port), RANK=str(rank), WORLD_SIZE=str(world_size))
        torch.set_num_threads(1)
        dist.init_process_group("gloo", rank=rank, world_size=world_size)
        args = ModelArgs(dim=32, n_layers=4, n_heads=4, vocab_size=61, ffn_dim=48, dropout=0.0)
        torch.manual_seed(5)
        full = Transformer(args)
        x, y = torch.randint(0, 61, (8, 8)), torch.randint(0, 61, (8, 8))
        oracle = copy.deepcopy(full)
        opt = torch.optim.AdamW(oracle.parameters(), lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.0)
        norms = []
        for _ in range(3):
            opt.zero_grad()
            torch.nn.functional.cross_entropy(oracle(x).reshape(-1, 61), y.reshape(-1)).backward()
            norms.append(float(torch.nn.utils.clip_grad_norm_(oracle.parameters(), max_norm)))
            opt.step()
        want = dict(oracle.state_dict())
        st = manual_model_split(copy.deepcopy(full), rank, world_size, torch.device("cpu"))
        sched = make_schedule("1F1B", st, 4, None,
                              optimizer={"lr": 1e-2, "betas": (0.9, 0.95), "eps": 1e-8, "weight_decay": 0.0, "max_grad_norm": max_norm})
        got_norms = []
        for _ in range(3):
            sched.step(*((x,) if rank == 0 else ()), target=y if rank == world_size - 1 else None,
                       losses=[] if rank == world_size - 1 else None)
            got_norms.append(float(sched.last_grad_norm_sq.sqrt()) / 4)        # grad_scale = 1/m
        err = max((v - want[k]).abs().max().item() for k, v in st.state_dict().items())
        out = {"rank": rank, "max_param_err": err, "norm_err": max(abs(a - b) for a, b in zip(norms, got_norms)),
               "clipped": norms[0] > max_norm}
        gathered = [None] * world_size
        dist.all_gather_object(gathered, out)
        if rank == 0:
            results_queue.put({"ranks": gathered})
        dist.barrier()
        dist.destroy_process_group()
    except Exception as e:  # noqa: BLE001
        traceback.print_exc()
        results_queue.put({"error": f"rank {rank}: {e}"})


def train_cli_worker(rank, world_size, argv, results_queue, master_port=None):
    """Run the training entry point (ppb200.train.main) as one rank of a gloo job."""
    try:
        import torch
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(master_port), RANK=str(rank), LOCAL_RANK=str(rank),
                          WORLD_SIZE=str(world_size))
        torch.set_num_threads(1)
        from ppb200 import train
        out = train.main(list(argv) + ["--device", "cpu"])
        if rank == world_size - 1:
            results_queue.put(out)
    except Exception as e:  # noqa: BLE001
        traceback.print_exc()
        results_queue.put({"error": f"rank {rank}: {e}"})


def tp_clip_worker(rank, world_size, results_queue, master_port=None, max_norm=0.05):
    """tp=2 (x pp): the global gradient norm counts sharded parameters once per shard and replicated ones once; the
    clipped AdamW update equals torch's on the unsharded model."""
    try:
        import torch
        import torch.distributed as dist
        from ppb200.models import ModelArgs, Transformer
        from ppb200.models.tp import shard_model_for_tp, shard_tensor
        from ppb200.api import make_schedule
        from ppb200.parallel.partition import prune_to_stage
        from ppb200.parallel.stage import PipelineStage
        from ppb200.parallel.mesh import init_mesh
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(master_port), RANK=str(rank), WORLD_SIZE=str(world_size))
        torch.set_num_threads(1)
        dist.init_process_group("gloo", rank=rank, world_size=world_size)
        mesh = init_mesh(dp_size=1, pp_size=world_size // 2, tp_size=2)
        P = mesh.pp_size
        args = ModelArgs(dim=64, n_layers=2 * P, n_heads=4, vocab_size=61, ffn_dim=96, dropout=0.0, causal=True,
                         cross_attn=False, activation="swiglu", norm_first=True, norm="rmsnorm", rope=True)
        torch.manual_seed(6)
        full = Transformer(args)
        x, y = torch.randint(0, 61, (8, 16)), torch.randint(0, 61, (8, 16))
        oracle = copy.deepcopy(full)
        opt = torch.optim.AdamW(oracle.parameters(), lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.0)
        norms = []
        for _ in range(2):
            opt.zero_grad()
            torch.nn.functional.cross_entropy(oracle(x).reshape(-1, 61), y.reshape(-1)).backward()
            norms.append(float(torch.nn.utils.clip_grad_norm_(oracle.parameters(), max_norm)))
            opt.step()
        want = dict(oracle.state_dict())
        mdl = copy.deepcopy(full)
        prune_to_stage(mdl, mesh.pp_idx, P)
        shard_model_for_tp(mdl, mesh.tp_idx, 2)
        st = PipelineStage(mdl, mesh.pp_idx, P, torch.device("cpu"), tp_group=mesh.tp_group)
        sched = make_schedule("1F1B", st, 4, None, pp_group=mesh.pp_group, pp_group_bwd=mesh.pp_group_bwd,
                              optimizer={"lr": 1e-2, "betas": (0.9, 0.95), "eps": 1e-8, "weight_decay": 0.0, "max_grad_norm": max_norm})
        first, last = 0 in sched.local, (P - 1) in sched.local
        got = []
        for _ in range(2):
            sched.step(*((x,) if first else ()), target=y if last else None, losses=[] if last else None)
            got.append(float(sched.last_grad_norm_sq.sqrt()) / 4)
        err = max((v - shard_tensor(k, want[k], args.n_heads, mesh.tp_idx, 2, True)).abs().max().item()
                  for k, v in st.state_dict().items())
        out = {"rank": rank, "max_param_err": err, "norm_err": max(abs(a - b) for a, b in zip(norms, got))}
        gathered = [None] * world_size
        dist.all_gather_object(gathered, out)
        if rank == 0:
            results_queue.put({"ranks": gathered})
        dist.barrier()
        dist.destroy_process_group()
    except Exception as e:  # noqa: BLE001
        traceback.print_exc()
        results_queue.put({"error": f"rank {rank}: {e}"})


def tp_dropout_worker(rank, world_size, results_queue, master_port=None, postln=False):
    """tp=2 with dropout on: the explicit backward is the gradient of the (fixed-mask) forward -- central finite
    difference along a random direction over all sharded and replicated parameters -- and the run is reproducible."""
    try:
        import torch
        import torch.distributed as dist
        from ppb200.models import ModelArgs, Transformer
        from ppb200.models.tp import shard_model_for_tp
        from ppb200.api import make_schedule
        from ppb200.parallel.stage import PipelineStage
        from ppb200.parallel.mesh import init_mesh
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(master_port), RANK=str(rank), WORLD_SIZE=str(world_size))
        torch.set_num_threads(1)
        dist.init_process_group("gloo", rank=rank, world_size=world_size)
        mesh = init_mesh(dp_size=1, pp_size=1, tp_size=2)
        args = ModelArgs(dim=64, n_layers=2, n_heads=4, vocab_size=61, ffn_dim=96, dropout=0.2, causal=True, cross_attn=False,
                         activation="gelu", norm_first=True, norm="layernorm")
        if postln:   # reference block with ReLU + hidden dropout
            args = ModelArgs(dim=64, n_layers=2, n_heads=4, vocab_size=61, ffn_dim=96, dropout=0.2)
        torch.manual_seed(8)
        full = Transformer(args).double()
        x, y = torch.randint(0, 61, (4, 16)), torch.randint(0, 61, (4, 16))

        def build():
            mdl = copy.deepcopy(full)
            shard_model_for_tp(mdl, mesh.tp_idx, 2)
            st = PipelineStage(mdl, 0, 1, torch.device("cpu"), tp_group=mesh.tp_group)
            st.materialize(dtype=torch.float32)
            return st, make_schedule("1F1B", st, 2, None, pp_group=mesh.pp_group)

        def loss_of(sched):
            losses = []
            sched.step(x, target=y, losses=losses)
            return float(torch.stack([l.reshape(()) for l in losses]).mean())

        st, sched = build()
        l0 = loss_of(sched)
        g = st.grad[:st.n_total].clone()
        st2, sched2 = build()
        same = loss_of(sched2) == l0 and torch.equal(st2.grad[:st2.n_total], g)
        # direction: replicated parameters get the same numbers on both ranks, shards get rank-specific ones
        v = torch.zeros_like(g)
        rep = set(st.tp_replicated)
        for n, o, shp in st.param_names:
            k = int(torch.Size(shp).numel())
            import zlib
            gen = torch.Generator().manual_seed(zlib.crc32(n.encode()) % 100000 + (0 if (o, k) in rep else 7 * (rank + 1)))
            v[o:o + k] = torch.randn(k, generator=gen)
        dot_sh = torch.zeros(1, dtype=torch.float64)
        dot_rep = torch.zeros(1, dtype=torch.float64)
        for n, o, shp in st.param_names:
            k = int(torch.Size(shp).numel())
            d = (g[o:o + k].double() * v[o:o + k].double()).sum()
            if (o, k) in rep:
                dot_rep += d
            else:
                dot_sh += d
        dist.all_reduce(dot_sh, group=mesh.tp_group)
        pred = float(dot_sh + dot_rep)
        eps = 1e-3
        vals = []
        for sgn in (+1.0, -1.0):
            stp, schp = build()
            stp.param_arena[:stp.n_total].add_(v, alpha=sgn * eps)
            vals.append(loss_of(schp))
        fd = (vals[0] - vals[1]) / (2 * eps)
        out = {"rank": rank, "reproducible": bool(same), "fd": fd, "pred": pred, "dropout_active": abs(l0) > 0}
        gathered = [None] * world_size
        dist.all_gather_object(gathered, out)
        if rank == 0:
            results_queue.put({"ranks": gathered})
        dist.barrier()
        dist.destroy_process_group()
    except Exception as e:  # noqa: BLE001
        traceback.print_exc()
        results_queue.put({"error": f"rank {rank}: {e}"})
