"""CPU test of the N>1 host logic: two gloo ranks each own the slice b2c_partition gives
them, produce their share of the batch (here with the CPU oracle standing in for the GPU kernel),
and all-reduce the order-independent checksum words; the reduced checksum must equal the one of the
whole batch computed in one process.  No data-path collective exists -- only this verification one."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import _oracle

N, HS, WS, HK, WK, MODE = 13, 12, 9, 4, 3, 2   # odd batch: ranks get 7 and 6 images


def _checksum_words(a):
    """numpy mirror of b2c_checksum_device: (wrapping sum, xor) of the IEEE-754 bit patterns."""
    bits = np.ascontiguousarray(a, dtype=np.float64).view(np.uint64).ravel()
    with np.errstate(over="ignore"):
        s = int(np.add.reduce(bits, dtype=np.uint64)) if bits.size else 0
    x = int(np.bitwise_xor.reduce(bits)) if bits.size else 0
    return s, x


def _batch():
    rng = np.random.default_rng(77)
    return rng.random((N, HS, WS)), rng.random((HK, WK))


def _to_i64(u):
    return u - (1 << 64) if u >= (1 << 63) else u


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import fftconvolution_b200 as b2c
    oracle = _oracle.load_oracle()
    src, ker = _batch()
    b, e = b2c.partition(N, world, rank)
    mine = np.stack([oracle.convolve_fftw(MODE, src[i], ker) for i in range(b, e)]) if e > b else np.zeros((0, HS, WS))
    s, x = _checksum_words(mine)
    t_sum = torch.tensor([_to_i64(s), e - b], dtype=torch.int64)
    dist.all_reduce(t_sum, op=dist.ReduceOp.SUM)          # int64 wraps modulo 2^64, like the device sum
    t_xor = torch.tensor([_to_i64(x)], dtype=torch.int64)
    dist.all_reduce(t_xor, op=dist.ReduceOp.BXOR)
    if rank == 0:
        np.save(os.path.join(out_dir, "reduced.npy"), np.array([t_sum[0].item(), t_sum[1].item(), t_xor[0].item()], dtype=np.int64))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_partition_and_checksum_reduce(tmp_path, oracle):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    got = np.load(os.path.join(tmp_path, "reduced.npy"))
    src, ker = _batch()
    whole = np.stack([oracle.convolve_fftw(MODE, src[i], ker) for i in range(N)])
    s, x = _checksum_words(whole)
    assert int(got[1]) == N
    assert int(got[0]) == _to_i64(s)
    assert int(got[2]) == _to_i64(x)
