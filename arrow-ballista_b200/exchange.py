"""Intra-node exchange of hash-partitioned batches (bgpu_all_to_all): the B200-native replacement, inside one
NVSwitch node, for the reference's shuffle files + Flight `do_get` + `ShuffleReaderExec` fan-in
(/root/reference/ballista/core/src/execution_plans/shuffle_writer.rs:351-385,
/root/reference/ballista/executor/src/flight_service.rs:124-166,
/root/reference/ballista/core/src/execution_plans/shuffle_reader.rs:539-624).

One process per GPU. Output partition p is owned by rank p % world. `torch.distributed` is only the rendezvous
(it carries the 128-byte ncclUniqueId); the data moves through libbgpu's own NCCL communicator on the ctx stream.
"""
import ctypes as C

from ._lib import lib
from .runtime import DeviceBatch


def owner_of(partition, world):
    """Rank that owns output partition `partition`."""
    return partition % world


def exchange_plan(all_offsets, rank, world, n_partitions):
    """Layout of the batch rank `rank` receives (bgpu_exchange_plan; pure host logic, no device).

    all_offsets[s] = the n_partitions+1 partition offsets of rank s's partitioned batch. Returns (recv_off, owned):
    recv_off[k * world + s] = first row of owned partition owned[k] coming from source s; recv_off[-1] = total rows.
    """
    flat = (C.c_uint64 * (world * (n_partitions + 1)))(*[int(x) for offs in all_offsets for x in offs])
    n_owned = len(range(rank, n_partitions, world))
    recv = (C.c_uint64 * (n_owned * world + 1))()
    got = C.c_uint32()
    total = lib().bgpu_exchange_plan(flat, rank, world, n_partitions, recv, C.byref(got))
    assert got.value == n_owned and total == recv[n_owned * world]
    return list(recv), list(range(rank, n_partitions, world))


class Exchange:
    """All-to-all of partitioned DeviceBatches between the ranks of one node."""

    def __init__(self, ctx, rank, world):
        # torch bundles its own libnccl.so.2: load it first so that libbgpu's dlopen("libnccl.so.2") reuses that copy
        # instead of pulling the system one into the process under the same soname
        import torch  # noqa: F401

        self.ctx, self.rank, self.world = ctx, int(rank), int(world)
        ident = [None]
        if self.rank == 0:
            buf = (C.c_char * 128)()
            ctx.check(lib().bgpu_nccl_unique_id(ctx._h, buf))
            ident[0] = bytes(buf)
        if self.world > 1:
            import torch.distributed as dist

            dist.broadcast_object_list(ident, src=0)
        self._comm = C.c_void_p()
        ctx.check(lib().bgpu_nccl_comm_create(ctx._h, ident[0], self.rank, self.world, C.byref(self._comm)))
        self.last_path, self.last_sent_bytes = "", None

    def all_to_all(self, partitioned, offsets, n_partitions):
        """partitioned / offsets: the result of BatchPartitioner.partition_device with the same n on every rank.
        Returns this rank's partitions from all ranks, grouped by (partition, source rank)."""
        offs = (C.c_uint64 * (n_partitions + 1))(*[int(o) for o in offsets])
        out = C.c_void_p()
        self.ctx.check(lib().bgpu_all_to_all(self.ctx._h, self._comm, self.rank, self.world, partitioned._h, offs,
                                             n_partitions, C.byref(out)))
        return DeviceBatch(self.ctx, out)

    def small_exchange(self, batch, key_cols, n_partitions):
        """bgpu_small_exchange: the exchange of a batch of a few rows as ONE ncclAllGather. COLLECTIVE. Returns the
        received DeviceBatch, or None when some rank's batch does not fit the small path (every rank gets None together
        and takes `partition_exchange` / `all_to_all`)."""
        from ._lib import BGPU_EUNSUPPORTED

        keys = (C.c_int32 * len(key_cols))(*key_cols)
        out = C.c_void_p()
        rc = lib().bgpu_small_exchange(self.ctx._h, self._comm, self.rank, self.world, batch._h, keys, len(keys),
                                       n_partitions, C.byref(out))
        if rc == BGPU_EUNSUPPORTED:
            return None
        self.ctx.check(rc)
        self.last_path, self.last_sent_bytes = "small_allgather", 0
        return DeviceBatch(self.ctx, out)

    def exchange_state(self, batch, key_cols, n_partitions, window_bytes=64 << 20):
        """Hash repartition + exchange of an aggregate state, cheapest path first: one all-gather for a few rows, the fused
        scatter into peer windows for null-free fixed-width states, hash partition + NCCL all-to-all otherwise. COLLECTIVE
        (the ranks take the same path: each step's refusal is decided from data every rank has)."""
        got = self.small_exchange(batch, key_cols, n_partitions)
        if got is not None:
            return got
        return self.partition_exchange(batch, key_cols, n_partitions, window_bytes=window_bytes)

    def partition_exchange(self, batch, key_cols, n_partitions, window_bytes=64 << 20):
        """Fused hash repartition + exchange (bgpu_partition_exchange): == all_to_all(*partition_device(batch)), with the
        scatter kernel storing straight into the owner ranks' receive windows over NVLink. COLLECTIVE. Batches the fused
        kernel does not take (nulls, utf8, ...) go through hash partition + NCCL all-to-all on every rank alike.
        Returns the received DeviceBatch; `last_sent_bytes` / `last_path` describe the call."""
        from ._lib import BGPU_EUNSUPPORTED
        from .partitioner import BatchPartitioner, Partitioning

        if getattr(self, "_peer", None) is None:
            self._peer = C.c_void_p()
            self.ctx.check(lib().bgpu_peer_exchange_create(self.ctx._h, self._comm, self.rank, self.world, window_bytes,
                                                           C.byref(self._peer)))
        keys = (C.c_int32 * len(key_cols))(*key_cols)
        out, sent = C.c_void_p(), C.c_uint64()
        rc = lib().bgpu_partition_exchange(self._peer, batch._h, keys, len(keys), n_partitions, C.byref(out), None, C.byref(sent))
        if rc == BGPU_EUNSUPPORTED:  # decided collectively inside the call: every rank lands here together
            part, offsets = BatchPartitioner.try_new(Partitioning.Hash(list(key_cols), n_partitions), self.ctx).partition_device(batch)
            self.last_path, self.last_sent_bytes = "nccl", None
            return self.all_to_all(part, offsets, n_partitions) if self.world > 1 else part
        self.ctx.check(rc)
        self.last_path, self.last_sent_bytes = "fused_peer", int(sent.value)
        return DeviceBatch(self.ctx, out)

    def close(self):
        if getattr(self, "_peer", None):
            lib().bgpu_peer_exchange_destroy(self._peer)
            self._peer = None
        if getattr(self, "_comm", None):
            lib().bgpu_nccl_comm_destroy(self._comm)
            self._comm = None

    def __del__(self):
        self.close()
