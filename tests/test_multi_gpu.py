"""Multi-GPU checks (need >= 2 CUDA devices; skipped otherwise): the NVLink symmetric-memory collectives and the
fused AdamW + parameter-push kernel must reproduce the NCCL path."""
import os
import socket
import traceback

import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, comm_kind: str, queue) -> None:
    try:
        flags = comm_kind.split('+')[1:]
        overlap, reshard, direct = 'overlap' in flags, 'reshard' in flags, 'direct' in flags
        comm_kind = comm_kind.split('+')[0]
        os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR='127.0.0.1',
                          MASTER_PORT=str(port), LLMT_COMM=comm_kind,
                          LLMT_MULTIMEM_RS='0' if 'pullrs' in flags else '1',
                          # the bit-for-bit comparisons below are about engine schedules: keep the attention backward's
                          # fp32 adds in a fixed order (kv blocks are otherwise shared between CTAs on this tiny model)
                          FLASH_ATTENTION_DETERMINISTIC='1')
        import torch.distributed as dist
        torch.cuda.set_device(rank)
        dist.init_process_group('nccl', rank=rank, world_size=world, device_id=torch.device('cuda', rank))
        from llm_training_b200.models import Llama, LlamaConfig
        from llm_training_b200.ops import fused_linear_cross_entropy, shift_labels
        from llm_training_b200.optim import ShardedOptimizer
        from llm_training_b200.parallel import apply_fully_sharded, init_mesh
        mesh = init_mesh(world, 1)
        torch.manual_seed(0)
        cfg = LlamaConfig(vocab_size=1024, hidden_size=256, intermediate_size=512, num_hidden_layers=2,
                          num_attention_heads=4, num_key_value_heads=2)
        with torch.device('cuda'):
            model = Llama(cfg).to(torch.float32 if 'fp32' in flags else torch.bfloat16)
        engine = apply_fully_sharded(model, mesh.dp, device=torch.device('cuda', rank), reshard_after_forward=reshard)
        assert engine.nvlink == (comm_kind == 'nvlink') and engine.reshard == reshard
        assert (engine.shard_arena is not None) == (reshard and comm_kind == 'nvlink')
        engine.overlap_optimizer = overlap          # optimizer of step k under the forward of step k+1
        engine.enable_direct_grad_accumulation(direct)   # wgrad GEMMs accumulate straight into the flat grads
        opt = ShardedOptimizer(torch.optim.AdamW(model.parameters(), lr=1e-3, weight_decay=0.1), [engine],
                               max_grad_norm=1.0)
        losses, norms = [], []
        for step in range(4):
            g = torch.Generator().manual_seed(1000 + step * world + rank)
            ids = torch.randint(0, 1024, (2, 256), generator=g).cuda()
            opt.zero_grad()
            out = model(input_ids=ids, skip_logits=True)
            loss = fused_linear_cross_entropy(out.normed_hidden_states, model.lm_head.weight,
                                              shift_labels(ids).flatten())
            loss.backward()
            opt.step()
            losses.append(float(loss))
            norms.append(float(opt.last_grad_norm))
        opt.synchronize()
        torch.cuda.synchronize()
        if reshard:
            assert all(u.is_sharded and not u.grad_live for u in engine.units)
        if comm_kind == 'nvlink':
            # gradients travel through the in-tree kernels in BOTH sharding modes (ZeRO-3: staging ring)
            assert engine.nvlink_grads, 'gradient reduce-scatter fell back to NCCL'
            want = 'pull' if ('pullrs' in flags or 'fp32' in flags or not engine.comm._scalars.multicast_ptr) \
                else 'multimem'
            assert engine.comm.last_rs_kind == want, (engine.comm.last_rs_kind, want)
            if reshard:
                assert sum(len(u.lazy_heads) for u in engine.units) == 2 * cfg.num_hidden_layers   # q|k|v + gate|up
        with engine.unsharded():
            checksum = [float(p.float().sum()) for p in model.parameters()]
        queue.put((rank, 'ok', (losses, norms, checksum)))
        dist.barrier()
        dist.destroy_process_group()
    except Exception:  # noqa: BLE001
        queue.put((rank, 'error', traceback.format_exc()))


def _run(comm_kind: str, world: int = 2, attempts: int = 3, worker=None, timeout: int = 240):
    import torch.multiprocessing as mp
    worker = worker or _worker
    ctx = mp.get_context('spawn')
    for attempt in range(attempts):
        queue = ctx.Queue()
        port = _free_port()
        procs = [ctx.Process(target=worker, args=(r, world, port, comm_kind, queue)) for r in range(world)]
        for p in procs:
            p.start()
        out, failure = {}, None
        for _ in range(world):
            rank, status, payload = queue.get(timeout=timeout)
            if status != 'ok':
                failure = (rank, payload)
                break
            out[rank] = payload
        if failure is None:
            for p in procs:
                p.join(timeout=60)
            return out
        for p in procs:                      # a rank failed: stop the others before deciding what to do
            p.kill()
            p.join(timeout=30)
        # the probed port can be taken between the probe and the rendezvous: that is not the code under test
        if 'EADDRINUSE' in failure[1] and attempt + 1 < attempts:
            continue
        raise AssertionError(f'rank {failure[0]} ({comm_kind}):\n{failure[1]}')


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs 2 GPUs')
def test_nvlink_engine_matches_nccl_engine():
    ref = _run('torch')
    got = _run('nvlink')
    for rank in (0, 1):
        l_ref, n_ref, c_ref = ref[rank]
        l_got, n_got, c_got = got[rank]
        for a, b in zip(l_ref, l_got):
            assert abs(a - b) < 2e-2, (l_ref, l_got)
        for a, b in zip(n_ref, n_got):
            assert abs(a - b) < 2e-2 * max(1.0, a), (n_ref, n_got)
        for a, b in zip(c_ref, c_got):
            assert abs(a - b) <= 2e-2 * max(1.0, abs(a)), (a, b)
    # every rank ends with the same parameters
    assert all(abs(a - b) < 1e-6 * max(1.0, abs(a)) for a, b in zip(got[0][2], got[1][2]))
    # the overlapped optimizer step (per-unit update + push + barrier on the side stream, first use waits on the
    # unit's event) must not change a single bit, on either backend
    for kind, base in (('nvlink+overlap', got), ('torch+overlap', ref)):
        ov = _run(kind)
        for rank in (0, 1):
            assert ov[rank][0] == base[rank][0], (kind, ov[rank][0], base[rank][0])
            assert ov[rank][2] == base[rank][2], kind
    # direct wgrad accumulation into the flat gradient buffers: every parameter's gradient must be in the buffer
    # BEFORE the unit's reduce-scatter starts (the accumulate order differs, so compare with a tolerance)
    for kind, base in (('nvlink+direct', got), ('torch+direct', ref)):
        dr = _run(kind)
        for rank in (0, 1):
            for a, b in zip(base[rank][0], dr[rank][0]):
                assert abs(a - b) < 5e-3, (kind, base[rank][0], dr[rank][0])
            for a, b in zip(base[rank][1], dr[rank][1]):
                assert abs(a - b) < 5e-3 * max(1.0, a), (kind, base[rank][1], dr[rank][1])
            for a, b in zip(base[rank][2], dr[rank][2]):
                assert abs(a - b) <= 5e-3 * max(1.0, abs(a)), (kind, a, b)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs 2 GPUs')
def test_nvlink_engine_exchanges_parameters_that_are_not_bf16():
    """fp32 parameters on the NVLink engine (the 32-true / *-mixed layouts): the update kernel cannot push them, so
    the step must fall back to exchanging the slices — every rank ends with the same weights, equal to the NCCL
    engine's (round-1 advisor finding: the slices were never exchanged and the ranks diverged from step 1)."""
    ref = _run('torch+fp32')
    got = _run('nvlink+fp32')
    assert all(abs(a - b) <= 1e-6 * max(1.0, abs(a)) for a, b in zip(got[0][2], got[1][2])), 'ranks diverged'
    for rank in (0, 1):
        for a, b in zip(ref[rank][0], got[rank][0]):
            assert abs(a - b) < 1e-3, (ref[rank][0], got[rank][0])
        for a, b in zip(ref[rank][2], got[rank][2]):
            assert abs(a - b) <= 1e-3 * max(1.0, abs(a)), (a, b)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs 2 GPUs')
def test_zero3_resharding_engine_matches_resident_engine():
    """ZeRO-3 mode (persistent slices only; full buffers gathered before a unit's use and dropped after it) with the
    NCCL gather and with the NVLink pull from the symmetric shard arena, with and without the overlapped
    optimizer, against the resident NCCL engine."""
    ref = {'torch': _run('torch'), 'nvlink': _run('nvlink'), 'nvlink+pullrs': _run('nvlink+pullrs')}
    # same reduce-scatter kernel on both sides of each comparison, so the losses must agree bit for bit: ZeRO-3 only
    # changes WHEN buffers exist, and on NVLink where the gradients are staged (symmetric ring vs resident arena)
    for kind, base in (('torch+reshard', 'torch'), ('nvlink+reshard', 'nvlink'),
                       ('nvlink+reshard+pullrs', 'nvlink+pullrs'), ('nvlink+reshard+overlap+direct', 'nvlink')):
        got = _run(kind)
        for rank in (0, 1):
            if 'direct' not in kind:
                assert got[rank][0] == ref[base][rank][0], (kind, got[rank][0], ref[base][rank][0])
            for a, b in zip(ref[base][rank][0], got[rank][0]):
                assert abs(a - b) < 5e-3, (kind, ref[base][rank][0], got[rank][0])
            for a, b in zip(ref[base][rank][2], got[rank][2]):
                assert abs(a - b) <= 5e-3 * max(1.0, abs(a)), (kind, a, b)
    # in-switch reduction vs peer pulls vs NCCL: different summation orders, same gradients within bf16 rounding
    for rank in (0, 1):
        for other in ('nvlink', 'nvlink+pullrs'):
            for a, b in zip(ref['torch'][rank][1], ref[other][rank][1]):
                assert abs(a - b) < 2e-2 * max(1.0, a), (other, ref['torch'][rank][1], ref[other][rank][1])


def _tp_worker(rank: int, world: int, port: int, comm_kind: str, queue) -> None:
    try:
        os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR='127.0.0.1',
                          MASTER_PORT=str(port), LLMT_COMM=comm_kind)
        import torch.distributed as dist
        torch.cuda.set_device(rank)
        dist.init_process_group('nccl', rank=rank, world_size=world, device_id=torch.device('cuda', rank))
        from llm_training_b200.models import Phi3, Phi3Config
        from llm_training_b200.ops import shift_labels
        from llm_training_b200.parallel import init_mesh
        from llm_training_b200.parallel.tensor_parallel import _FusedGatherLinear
        mesh = init_mesh(1, world)
        torch.manual_seed(0)
        cfg = Phi3Config(vocab_size=1024, hidden_size=256, intermediate_size=512, num_hidden_layers=2,
                         num_attention_heads=4, num_key_value_heads=2, pad_token_id=0)
        with torch.device('cuda'):
            model = Phi3(cfg).to(torch.bfloat16)
        model.parallelize(dp_mesh=mesh.dp, tp_mesh=mesh.tp)
        tp = model.tp_context
        assert (tp.comm is not None) == (comm_kind == 'nvlink')
        ids = torch.randint(0, 1024, (2, 256), generator=torch.Generator().manual_seed(5)).cuda()
        out = model(input_ids=ids, skip_logits=True)
        loss = tp.vocab_parallel_cross_entropy(out.normed_hidden_states, model.lm_head.weight,
                                               shift_labels(ids).flatten())
        loss.backward()
        grads = [float(p.grad.float().abs().sum()) for p in model.parameters()]
        extra = None
        if comm_kind == 'nvlink':
            # vocabulary-parallel embedding as a scatter over NVLink: bit-equal to a lookup in the whole table
            shard = model.embed_tokens.weight.detach()
            table = torch.empty(shard.shape[0] * world, shard.shape[1], dtype=shard.dtype, device='cuda')
            dist.all_gather_into_tensor(table, shard.contiguous())
            want = table[ids.flatten()]
            want[ids.flatten() == cfg.pad_token_id] = 0
            for _ in range(3):          # repeated use walks the symmetric buffer ring
                got = tp.vocab_parallel_embedding(ids, model.embed_tokens.weight, cfg.pad_token_id, None, 0)
            assert torch.equal(got, want.chunk(world)[rank]), 'vocab-parallel embedding differs from the table lookup'
            # loss statistics combined by the NVLink kernel against the NCCL + torch combination
            from llm_training_b200.parallel.tensor_parallel import _VocabParallelLoss
            labels = shift_labels(ids).flatten()
            h = out.normed_hidden_states.detach()
            nll_fused = _VocabParallelLoss.apply(h, model.lm_head.weight.detach(), labels, -100, tp.axis, tp)
            nll_nccl = _VocabParallelLoss.apply(tp.gather_sequence(h), model.lm_head.weight.detach(), labels, -100,
                                                tp.axis)
            assert float((nll_fused - nll_nccl).abs().max()) < 1e-4, float((nll_fused - nll_nccl).abs().max())
            assert float(nll_fused[labels == -100].abs().max()) == 0.0
            # the fused all-gather -> GEMM kernel against NCCL all-gather + GEMM
            x = torch.randn(256, 256, device='cuda', generator=torch.Generator('cuda').manual_seed(rank)).bfloat16()
            w = (torch.randn(384, 256, device='cuda', generator=torch.Generator('cuda').manual_seed(7)) * 0.05).bfloat16()
            y = _FusedGatherLinear.apply(x, tp, w)
            full = torch.empty(256 * world, 256, dtype=x.dtype, device='cuda')
            dist.all_gather_into_tensor(full, x)
            extra = float((y.float() - full.float() @ w.float().t()).abs().max())
            # land-once gather GEMM: several row blocks per rank, both CTA-group forms, both B layouts; the
            # gathered operand it leaves behind must be exact
            from llm_training_b200.parallel.tensor_parallel import _gather_gemm
            for rows, k, n, b_mn in ((1024, 512, 768, False), (384, 256, 320, False), (512, 384, 512, True)):
                gen = torch.Generator('cuda').manual_seed(100 + rank)
                xl = torch.randn(rows, k, device='cuda', generator=gen).bfloat16()
                gw = torch.Generator('cuda').manual_seed(9)
                wl = (torch.randn((k, n) if b_mn else (n, k), device='cuda', generator=gw) * 0.05).bfloat16()
                want_full = torch.empty(rows * world, k, dtype=xl.dtype, device='cuda')
                dist.all_gather_into_tensor(want_full, xl)
                for _ in range(3):      # repeated use exercises the flag reset and the buffer ring
                    got, got_full = _gather_gemm(tp, xl, wl, b_mn)
                want = want_full.float() @ (wl.float() if b_mn else wl.float().t())
                assert torch.equal(got_full, want_full), 'gathered operand differs'
                extra = max(extra, float((got.float() - want).abs().max()))
        torch.cuda.synchronize()
        queue.put((rank, 'ok', (float(loss), grads, extra)))
        dist.barrier()
        dist.destroy_process_group()
    except Exception:  # noqa: BLE001
        queue.put((rank, 'error', traceback.format_exc()))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs 2 GPUs')
def test_fused_tensor_parallel_kernels_match_nccl_path():
    ref, got = _run('torch', worker=_tp_worker), _run('nvlink', worker=_tp_worker)
    for rank in (0, 1):
        assert abs(ref[rank][0] - got[rank][0]) < 2e-2, (ref[rank][0], got[rank][0])
        for a, b in zip(ref[rank][1], got[rank][1]):
            assert abs(a - b) <= 5e-2 * max(1e-3, abs(a)), (a, b)
        assert got[rank][2] is not None and got[rank][2] < 5e-2


def _flag_stress_worker(rank: int, world: int, port: int, comm_kind: str, queue) -> None:
    try:
        os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR='127.0.0.1',
                          MASTER_PORT=str(port), LLMT_COMM='nvlink')
        import torch.distributed as dist
        torch.cuda.set_device(rank)
        dist.init_process_group('nccl', rank=rank, world_size=world, device_id=torch.device('cuda', rank))
        from llm_training_b200.parallel import init_mesh
        from llm_training_b200.parallel.comm import get_comm
        mesh = init_mesh(world, 1)
        comm = get_comm(mesh.dp)
        assert getattr(comm, 'is_nvlink', False)
        n = 1 << 16
        slice_n = n // world
        buf = comm.alloc(n, torch.float32)
        bad = 0
        val = torch.zeros(4, dtype=torch.float32, device='cuda')
        for i in range(1500):
            # every rank stamps ITS slice with the round number (the rest with -1), then pulls the peers' slices
            # through the mapped pointers (barrier + pull kernel): a stale stamp means the flag barrier let a reader
            # through before the writer's data was visible, or let a writer of round i+1 overtake a reader of round i
            buf.tensor.fill_(-1.0)
            buf.tensor[rank * slice_n:(rank + 1) * slice_n].fill_(float(i))
            if (i + rank) % 3 == 0:
                torch.cuda._sleep(200_000)              # skew the ranks
            comm.allgather_pull(buf, 0, slice_n * 4)
            peer = (rank + 1) % world
            stamps = buf.tensor[peer * slice_n:(peer + 1) * slice_n]
            if float(stamps.min()) != float(i) or float(stamps.max()) != float(i):
                bad += 1
            comm.barrier()                              # nobody overwrites before everyone has read
            val.fill_(float(rank + i))
            comm.all_reduce(val)
            expect = sum(r + i for r in range(world))
            if abs(float(val[0]) - expect) > 1e-3:
                bad += 1
        torch.cuda.synchronize()
        queue.put((rank, 'ok', bad))
        dist.barrier()
        dist.destroy_process_group()
    except Exception:  # noqa: BLE001
        queue.put((rank, 'error', traceback.format_exc()))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs 2 GPUs')
def test_symmetric_memory_flag_protocol_stress():
    """1500 rounds of write -> flag barrier -> peer read -> barrier -> one-shot all-reduce with skewed ranks: the
    epoch flags must never let a reader run ahead of a writer."""
    out = _run('nvlink', worker=_flag_stress_worker, timeout=300)
    for rank, payload in out.items():
        assert payload == 0, f'rank {rank}: {payload} stale reads / wrong sums'
