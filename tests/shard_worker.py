"""One rank of the sample-sharded path, launched by torch.distributed.run (rendezvous on 127.0.0.1).

  nccl : one GPU per rank, ShardedCaller (the C ABI's NCCL exchange); rank 0 checks the gathered records against
         the per-read oracle on the WHOLE cohort.
  gloo : CPU only, no CUDA call -- the host-side protocol: identical tiles from the seed on every rank, sample-range
         sharding, the dense exchange layout (shard_ref.dense_hist), all-reduce of the integer histograms over gloo,
         site-slice ownership; the owner runs the ORACLE on the exchanged statistics (shard_ref.expand) and rank 0
         checks the gathered slices against the oracle on the whole cohort.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)


def main():
    mode = sys.argv[1]
    import torch
    import torch.distributed as dist
    import basevarc_b200 as B
    import oracle_py
    from parity import compare_records
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist.init_process_group("gloo")            # plumbing only: id broadcast + gathering the records for the check
    cases = [(20000, 257, 0.1, "highpoly", 4), (1000, 1500, 0.5, "lowpass", 2)]
    caller = None
    for ci, (N, S, depth, model, seed) in enumerate(cases):
        tile = B.synth_tile(N, S, depth, model, seed=seed)          # every rank derives the same cohort from the seed
        mine = tile.shard_samples(world)[rank]
        sr = (S + world - 1) // world
        if mode == "nccl":
            uid = [B.shard_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(uid, src=0)
            caller = B.ShardedCaller(N, rank, world, uid[0], device=local, max_sites=S, max_entries=max(mine.n_entries, 1))
            first, rec = caller.shard_call(mine)
            first2, rec2 = caller.shard_run()                        # resident re-run gives the same bytes
            assert first2 == first and rec2.tobytes() == rec.tobytes()
            st = caller.shard_stats()
            assert caller.kernel_launches > 0 and st["exchange_bytes"] > 0
            caller.close()
        else:
            import shard_ref
            qlo, qhi = 1, 63
            dense = torch.from_numpy(shard_ref.dense_hist(mine, qlo, qhi).astype(np.int64))
            dist.all_reduce(dense)                                    # gloo has no reduce-scatter: all-reduce, keep the slice
            first = rank * sr
            own = dense.numpy()[first:min(S, first + sr)].astype(np.uint32)
            sub = shard_ref.expand(own, tile.ref_base[first:first + own.shape[0]], N, qlo, qhi)
            rec = oracle_py.call_sites(sub, B.default_params(N))
        assert first == rank * sr
        parts = [None] * world
        dist.all_gather_object(parts, rec)
        if rank == 0:
            got = np.concatenate(parts)
            want = oracle_py.call_sites(tile, B.default_params(N))
            if mode == "gloo":       # mapq / rpr do not travel in the first exchange: the rank-sum fields are not comparable
                got = got.copy(); got["ranksum"] = want["ranksum"]
            nbad, lines = compare_records(got, want)
            assert nbad == 0, "\n".join(lines)
            assert want["lrt_ok"].sum() > 0
            print("case %d: %d sites, %d LRT-positive, 0 mismatches" % (ci, S, int(want["lrt_ok"].sum())), flush=True)
        if mode == "gloo" and ci == 0:
            # third exchange: the rank-sum histograms of an LRT-positive site are additive over the sample shards
            want = oracle_py.call_sites(tile, B.default_params(N))
            site = int(np.nonzero(want["lrt_ok"] == 1)[0][0])
            am = 0
            for b in want["alt_bases"][site][:want["n_alt"][site]]:
                am |= 1 << int(b)
            h = torch.from_numpy(shard_ref.vcf_hist(mine, site, int(tile.ref_base[site]), am).astype(np.int64))
            dist.all_reduce(h)
            assert np.array_equal(h.numpy(), shard_ref.vcf_hist(tile, site, int(tile.ref_base[site]), am).astype(np.int64))
    dist.barrier()
    if rank == 0:
        print("shard_worker ok", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
