"""CPU: DistributedFusedAdam checkpoint shards ({step, fp32_p, fp32_m, fp32_v} per rank, reference
apex/contrib/optimizers/distributed_fused_adam.py:598-636) re-sliced for another world size / chunk count:
pure layout arithmetic first, then a job of 2 ranks saved after two steps and resumed as a job of 1 rank,
which must continue bit-identically (kernel doubles call the oracle, as in test_dist_adam_cpu.py)."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "rocm-apex_b200"))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

SHAPES = [(300, 7), (64,), (5,), (1000, 3), (1,), (129,)]


def _shard(full, layout, rank):
    K, chunk, shard = layout["num_chunks"], layout["chunk"], layout["shard"]
    return torch.cat([full[k * chunk + rank * shard:k * chunk + (rank + 1) * shard] for k in range(K)]).clone()


def test_reshard_is_a_reslicing_of_the_flat_space():
    from apex.contrib.optimizers.distributed_fused_adam import build_layout, reshard_state_dicts
    numels = [300 * 7, 64, 5, 3000, 1, 129, 70000]
    g = torch.Generator().manual_seed(3)
    end = build_layout(numels, 1, 1)["offsets"][-1] + numels[-1]
    for (W0, K0), (W1, K1) in [((4, 3), (2, 5)), ((2, 1), (1, 1)), ((1, 4), (8, 2)), ((8, 16), (3, 1)), ((3, 2), (3, 2))]:
        old, new = build_layout(numels, W0, K0), build_layout(numels, W1, K1)
        fulls = {}
        for key in ("fp32_p", "fp32_m", "fp32_v"):
            f = torch.zeros(max(old["total"], new["total"]))
            f[:end] = torch.randn(end, generator=g)
            fulls[key] = f
        sds = [dict(step=17, **{k: _shard(f[:old["total"]], old, r) for k, f in fulls.items()}) for r in range(W0)]
        out = reshard_state_dicts(sds, numels, W1, K1, old_num_chunks=K0)
        assert len(out) == W1
        for r, sd in enumerate(out):
            assert sd["step"] == 17
            for k, f in fulls.items():
                assert torch.equal(sd[k], _shard(f[:new["total"]], new, r)), (W0, K0, W1, K1, r, k)
    # the chunk count is inferred when the shard length pins it, and refused when it does not
    sds = [dict(step=1, **{k: _shard(torch.zeros(build_layout(numels, 2, 3)["total"]), build_layout(numels, 2, 3), r)
                           for k in ("fp32_p", "fp32_m", "fp32_v")}) for r in range(2)]
    try:
        reshard_state_dicts(sds, numels, 1, 1)
    except ValueError as e:
        assert "old_num_chunks" in str(e)
    sds[1]["step"] = 2
    with pytest.raises(ValueError):
        reshard_state_dicts(sds, numels, 1, 1, old_num_chunks=3)


def test_reshard_tool_command_line(tmp_path):
    import json
    import subprocess
    from apex.contrib.optimizers.distributed_fused_adam import build_layout
    numels = [2100, 64, 5, 3000, 1, 129]
    old, new = build_layout(numels, 2, 2), build_layout(numels, 4, 1)
    g = torch.Generator().manual_seed(9)
    end = old["offsets"][-1] + numels[-1]
    fulls = {}
    for key in ("fp32_p", "fp32_m", "fp32_v"):
        f = torch.zeros(max(old["total"], new["total"]))
        f[:end] = torch.randn(end, generator=g)
        fulls[key] = f
    files = []
    for r in range(2):
        sd = dict(step=5, layout={"world_size": 2, "num_chunks": 2},
                  **{k: _shard(f[:old["total"]], old, r) for k, f in fulls.items()})
        files.append(str(tmp_path / ("old%d.pt" % r)))
        torch.save(sd, files[-1])
    (tmp_path / "numels.json").write_text(json.dumps(numels))
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "reshard_checkpoint.py"), "--numels", str(tmp_path / "numels.json"),
                          "--new-world", "4", "--new-chunks", "1", "--out", str(tmp_path / "new{rank}.pt")] + files,
                         capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    for r in range(4):
        sd = torch.load(str(tmp_path / ("new%d.pt" % r)))
        assert sd["step"] == 5 and sd["layout"] == {"world_size": 4, "num_chunks": 1}
        for k, f in fulls.items():
            assert torch.equal(sd[k], _shard(f[:new["total"]], new, r))


def _grads(it, rank):
    gr = torch.Generator().manual_seed(100 * it + rank)
    return [torch.randn(s, generator=gr) * 4.0 for s in SHAPES]


def _worker(rank, world, port, tmp, phase, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    for p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import apex  # noqa: F401
        from test_dist_adam_cpu import _install_kernel_doubles
        _install_kernel_doubles()
        from apex.contrib.optimizers import DistributedFusedAdam, reshard_state_dicts
        g0 = torch.Generator().manual_seed(7)
        params = [torch.nn.Parameter(torch.randn(s, generator=g0) * 0.1) for s in SHAPES]
        kw = dict(lr=1e-2, betas=(0.9, 0.99), eps=1e-8, weight_decay=0.01, dwu_num_chunks=1, overlap_reductions=False)
        if phase == "write":                      # 2 ranks, 2 pipeline chunks
            opt = DistributedFusedAdam(params, dwu_num_blocks=2, **kw)
            opt.set_global_scale(4.0)
            for it in range(3):
                for p, g in zip(params, _grads(it, rank)):
                    p.grad = g
                opt.step()
                if it == 1:
                    torch.save({k: (v.clone() if torch.is_tensor(v) else v) for k, v in opt.state_dict().items()},
                               os.path.join(tmp, "old_rank%d.pt" % rank))
            torch.save([p.detach().clone() for p in params], os.path.join(tmp, "final_rank%d.pt" % rank))
        else:                                     # 1 rank, 3 pipeline chunks, resumed from the re-sliced shards
            sds = [torch.load(os.path.join(tmp, "old_rank%d.pt" % r)) for r in range(2)]
            new = reshard_state_dicts(sds, [p.numel() for p in params], 1, 3)
            opt = DistributedFusedAdam(params, dwu_num_blocks=3, **kw)
            opt.load_state_dict(new[0])
            opt.set_global_scale(4.0)
            g0s, g1s = _grads(2, 0), _grads(2, 1)
            for p, a, b in zip(params, g0s, g1s):
                p.grad = a / 2 + b / 2            # what the 2-rank reduce-scatter of predivided gradients produced
            opt.step()
            want = torch.load(os.path.join(tmp, "final_rank0.pt"))
            for p, w in zip(params, want):
                assert torch.equal(p.detach(), w)
            assert opt.state_dict()["step"] == 3
        out.put((rank, "ok"))
    except Exception:
        import traceback
        out.put((rank, "FAIL: " + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_two_rank_checkpoint_resumes_on_one_rank(tmp_path):
    ctx = mp.get_context("spawn")
    for phase, world, port in (("write", 2, 29690), ("resume", 1, 29691)):
        q = ctx.Queue()
        procs = [ctx.Process(target=_worker, args=(r, world, port, str(tmp_path), phase, q)) for r in range(world)]
        for p in procs:
            p.start()
        results = [q.get(timeout=240) for _ in procs]
        for p in procs:
            p.join(timeout=60)
        for rank, msg in results:
            assert msg == "ok", "%s rank %d: %s" % (phase, rank, msg)
