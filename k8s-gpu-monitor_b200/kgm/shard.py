"""Multi-GPU plumbing: partition of the fleet by host and assembly of the ranks' outputs.

Every series key of the reference starts with ``host=`` (gpu.py:137,207,241) and every join
input is node-local, so whole hosts are the partition unit: all documents of a host go to
one rank and no sample ever crosses GPUs.  The only exchange is (1) an all-gather of each
rank's ``{bytes, lines}`` so that every rank knows where its slice of the fleet-wide body
starts, and (2) the concatenation of the encoded buffers, rank-major.  One process per GPU;
``torch.distributed`` (NCCL on the GPU box, gloo in the CPU tests) carries both.
"""
from __future__ import annotations

import hashlib

import os

import torch
import torch.distributed as dist


def host_hash(name: str) -> int:
    return int.from_bytes(hashlib.blake2b(name.encode(), digest_size=8).digest(), "little")


def assign_hosts(names, weights, world: int):
    """host -> rank.  Size-aware (largest first onto the least loaded rank) so that skewed
    hosts do not unbalance the shards; ties and order are fixed by a hash of the host name,
    so every rank computes the same map without talking."""
    order = sorted(range(len(names)), key=lambda i: (-int(weights[i]), host_hash(names[i]), names[i]))
    load = [0] * world
    out = [0] * len(names)
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        out[i] = r
        load[r] += int(weights[i])
    return out


def gather_counts(n_bytes: int, n_lines: int, device, group=None) -> torch.Tensor:
    """All-gather of the per-rank result sizes -> int64 [world, 2]."""
    world = dist.get_world_size(group)
    mine = torch.tensor([n_bytes, n_lines], dtype=torch.int64, device=device)
    out = torch.empty((world, 2), dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(out.view(-1), mine, group=group)
    return out


def slice_offsets(counts: torch.Tensor):
    """Exclusive scan of the gathered counts: where each rank's bytes / lines start."""
    c = counts.cpu()
    off = torch.zeros((c.shape[0] + 1, 2), dtype=torch.int64)
    off[1:] = torch.cumsum(c, 0)
    return off


def concat_encoded(local: torch.Tensor, counts: torch.Tensor, out: torch.Tensor = None, group=None):
    """Rank-major concatenation of every rank's encoded bytes on every rank.

    ``local`` is this rank's uint8 buffer (its first ``counts[rank,0]`` bytes are valid).
    The variable-length all-gather is one padded ``all_gather_into_tensor`` followed by a
    compaction of the padding on the device."""
    world = dist.get_world_size(group)
    off = slice_offsets(counts)
    sizes = counts[:, 0].cpu().tolist()
    maxb = max(max(sizes), 1)
    maxb = (maxb + 15) // 16 * 16
    padded = torch.empty(maxb, dtype=torch.uint8, device=local.device)
    padded[:sizes[dist.get_rank(group)]] = local[:sizes[dist.get_rank(group)]]
    gathered = torch.empty((world, maxb), dtype=torch.uint8, device=local.device)
    dist.all_gather_into_tensor(gathered.view(-1), padded, group=group)
    total = int(off[-1, 0])
    if out is None:
        out = torch.empty(max(total, 1), dtype=torch.uint8, device=local.device)
    for r in range(world):
        out[int(off[r, 0]):int(off[r, 0]) + sizes[r]] = gathered[r, :sizes[r]]
    return out[:total], off


class ShardGather:
    """Brings every step's encoded shards together on one GPU, rank-major.

    Three ways, chosen at construction:

    ``push`` (CUDA, ``push=True``, needs CUDA IPC between the ranks): the gathering GPU owns
    fixed-stride slots that every rank maps; the library's kernels write each shard and its
    {n_bytes, n_lines} header straight into its slot while they encode, and the ranks tell each
    other what has landed through counters in the same mapped memory (``push_step``: one
    library call per step, no collective).  With ``rotate_root`` step k is gathered on rank
    k mod world.

    ``slots`` (CUDA otherwise): the shards travel as fixed-size NCCL transfers into the
    fixed-stride slots -- ``start_device``: one all-gather straight out of the library's
    buffers, sizes in the headers, nothing known on the host; ``start``: the older per-step
    path with a local copy, a host wait and the packing kernel every step.

    ``p2p`` (gloo CPU tests, ``host_counts=True``): all-gather of {bytes, lines}, then every
    shard goes to its final offset on rank 0 as a send/recv sized on the host.

    In the slot modes the slots are the gathered result; ``pack()`` / ``body()`` close the
    gaps with the library's kgm_concat_shards kernel.  ``finish`` waits for every transfer."""

    def __init__(self, cap_bytes: int, device, group=None, host_counts=False, push=False,
                 rotate_root=False, contexts=1):
        self.group, self.device = group, device
        # push mode: how many library contexts the caller visits in turn (step k runs on context
        # k mod contexts, on every rank alike); every (rank, context) gets its own "landed" counter
        self.contexts = max(1, int(contexts))
        self.rotate_root = bool(rotate_root)
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.cap = int(cap_bytes)
        if self.world > 1:                           # every rank must size its buffers alike
            t = torch.tensor([self.cap], dtype=torch.int64, device=device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
            self.cap = int(t.item())
        self.xbuf = [torch.empty(self.cap, dtype=torch.uint8, device=device) for _ in range(2)]
        self.final = (torch.empty(self.cap * self.world, dtype=torch.uint8, device=device)
                      if self.rank == 0 else None)
        # the tiny count exchange gets its own communicator, so that it never queues behind the
        # bulk shard transfers of the previous step
        self.ctl_group = dist.new_group(backend=dist.get_backend(group)) if self.world > 1 else None
        self.pending = [[], []]
        self.k = 0
        self.last_offsets = None
        self.cuda = device.type == "cuda"
        self.stream = torch.cuda.Stream(device) if self.cuda else None
        self.copied = [torch.cuda.Event(), torch.cuda.Event()] if self.cuda else None
        self.mode = "p2p"
        if self.cuda and self.mode == "p2p" and self.world > 1 and not host_counts:
            self._init_slots()
            if push:
                self._open_push()

    # per gathering rank: landed[contexts][world] (u32; one counter per writing stream, so that a plain
    # store is as good as an atomic max), then ack (u32)
    @property
    def FLAG_BYTES(self):
        return (4 * (self.contexts * self.world + 1) + 255) // 256 * 256

    def _ack_idx(self):
        return self.contexts * self.world

    def _landed_idx(self, k, rank):
        return (k % self.contexts) * self.world + rank


    def _open_push(self):
        """Push mode: every gathering rank owns PUSH_SETS sets of fixed-stride slots plus a block
        of counters, and every other rank maps them (CUDA IPC, NVLink peer access).  A step on
        a rank is ONE library call (kgm_push_step) that queues, on the context's stream: wait
        for a counter, the poll kernel writing the shard and its size header straight into its
        slot on the step's gathering rank (kgm_set_output_buffer), and a store of "landed" to
        that rank's counter block.  The gathering rank of a step waits for all `world` landed
        counters on a stream of its own and then bumps its `ack` counter; writers read it to
        know how far everybody has come.  No collective runs per step.  With ``rotate_root``
        the gathering rank of step k is k mod world: every step's concatenation still lands
        whole on one GPU, but the ingress of (world - 1) shards per step is spread over all
        NVLink ports instead of one."""
        ok = torch.ones(1, dtype=torch.int32, device=self.device)
        # who carries a shard to the gathering GPU: a copy engine behind the poll kernel (the
        # kernel then writes local memory only and never waits for another GPU; default), or the
        # poll kernel's own stores into the peer slot (KGM_PUSH_KERNEL_STORES=1)
        self.copy_engine = os.environ.get("KGM_PUSH_KERNEL_STORES", "0") != "1"
        # How far a rank may run ahead of the last step that has landed everywhere.  A whole
        # step lands on ONE GPU (7 x 46 MB at N=8: 0.36 ms of its NVLink ingress), so several
        # steps, bound for different gathering GPUs, must be in flight at once.  Copy engines:
        # 2 x world steps for every rank (each context's copy stream carries one).  Kernel
        # stores with a rotating root: windows that differ by rank (rank r: world - r + 1 steps),
        # so that the ranks settle one step apart and write to `world` different GPUs.
        skew = self.rotate_root and os.environ.get("KGM_PUSH_SKEW", "1") == "1"
        if self.copy_engine and self.rotate_root:
            self.push_ahead = 2 * self.world
        else:
            self.push_ahead = (self.world - self.rank + 1) if skew else 3
        if os.environ.get("KGM_PUSH_AHEAD"):         # (experiments) at most (PUSH_SETS - 1) x world: a set must not be overtaken
            self.push_ahead = max(1, min(int(os.environ["KGM_PUSH_AHEAD"]), 3 * self.world if self.rotate_root else 3))
        self.PUSH_SETS = 4                           # a set is reused when its step is > world + 1 steps back
        self.flag_off = self.PUSH_SETS * self.world * self.stride
        total = self.flag_off + self.FLAG_BYTES
        roots = range(self.world) if self.rotate_root else (0,)
        self.push_bufs = {}
        try:
            mine = None
            if self.rank in roots:
                mine = torch.zeros(total, dtype=torch.uint8, device=self.device)
                self.push_bufs[self.rank] = mine
            boxes = [None] * self.world
            dist.all_gather_object(boxes, mine.untyped_storage()._share_cuda_() if mine is not None else None,
                                   group=self.group)
            for r in roots:
                if r == self.rank:
                    continue
                # open the handle on THIS rank's device (not on the exporter's index): the
                # mapping must belong to the context whose kernels will write through it
                handle = (self.device.index,) + tuple(boxes[r][1:])
                with torch.cuda.device(self.device):
                    storage = torch.UntypedStorage._new_shared_cuda(*handle)
                buf = torch.empty(0, dtype=torch.uint8, device=storage.device).set_(storage)
                if buf.numel() < total:
                    raise RuntimeError("mapped buffer is smaller than expected")
                self.push_bufs[r] = buf
        except Exception:
            ok.zero_()
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)
        if int(ok.item()) != 1:
            self.push_bufs = {}
            return
        torch.cuda.synchronize(self.device)          # the zeroed counters are in place before anybody writes
        dist.barrier(group=self.group)
        self.push_step_no = 0
        self.last_root = 0
        self.mode = "push"

    def push_root(self, k=None):
        """The rank that gathers step k (default: the next step)."""
        k = self.push_step_no if k is None else k
        return k % self.world if self.rotate_root else 0

    def _uses(self, k):
        return k // self.world if self.rotate_root else k          # how often step k's root has gathered before

    def _push_offset(self, k):
        return (self._uses(k) % self.PUSH_SETS) * self.world * self.stride

    def _flag_ptr(self, root, idx):
        return self.push_bufs[root].data_ptr() + self.flag_off + 4 * idx

    def push_step(self, mon):
        """One step: queues on ``mon``'s stream (kgm_push_step) the wait for the step
        ``push_ahead`` back to have landed everywhere, the encode of ``mon``'s resident batch
        into this rank's slot on the step's gathering rank, and the "landed" store.  Returns
        the library's (unsized) device result; finish with ``mon.encode_finish()``."""
        import ctypes as C
        k = self.push_step_no
        self.push_step_no += 1
        self.k += 1
        root, uses = self.push_root(k), self._uses(k)
        slot = self.push_bufs[root].data_ptr() + self._push_offset(k) + self.rank * self.stride
        j = k - self.push_ahead
        wait_flag, wait_value = (self._flag_ptr(self.push_root(j), self._ack_idx()), self._uses(j) + 1) if j >= 0 else (0, 0)
        res = mon.push_step(slot, self.stride, wait_flag, wait_value, self._flag_ptr(root, self._landed_idx(k, self.rank)), uses + 1,
                            copy_engine=self.copy_engine)
        self._packed = False
        self.last_root = root
        if self.rank == root:
            # this rank gathers the step: all shards landed -> acknowledge, on the gather's own stream
            st = C.c_void_p(self.stream.cuda_stream)
            rc = self._lib.kgm_stream_wait_flags(st, C.c_void_p(self._flag_ptr(root, self._landed_idx(k, 0))), self.world,
                                                 (uses + 1) & 0xFFFFFFFF)
            ack = (C.c_void_p * 1)(self._flag_ptr(root, self._ack_idx()))
            rc = rc or self._lib.kgm_stream_signal_flags(st, ack, 1, (uses + 1) & 0xFFFFFFFF)
            if rc != 0:
                raise RuntimeError("flag kernels failed: %d" % rc)
            base = self._push_offset(k)
            self.slots = self.push_bufs[self.rank][base:base + self.world * self.stride]
            self.all_slots = self.slots
            if self.final is None:
                self.final = torch.empty(self.cap * self.world, dtype=torch.uint8, device=self.device)
        return res

    def _init_slots(self):
        """CUDA, default mode: fixed-stride slots on rank 0, sizes stay on the device."""
        from . import _ffi
        self._lib = _ffi.load()
        self.stride = (self.cap + 16 + 15) // 16 * 16          # 16-byte size header + shard
        for i in range(2):
            self.xbuf[i] = torch.empty(self.stride, dtype=torch.uint8, device=self.device)
        self.mine_host = [torch.zeros(2, dtype=torch.int64).pin_memory() for _ in range(2)]
        self.mine_dev = [torch.zeros(2, dtype=torch.int64, device=self.device) for _ in range(2)]
        self.counts_dev = [torch.zeros((self.world, 2), dtype=torch.int64, device=self.device) for _ in range(2)]
        self.totals_dev = torch.zeros(2, dtype=torch.int64, device=self.device)
        if self.rank == 0:
            self.slots = torch.empty(self.stride * self.world, dtype=torch.uint8, device=self.device)
            self.own = [torch.empty(self.stride, dtype=torch.uint8, device=self.device) for _ in range(2)]
        self.copy_stream = torch.cuda.Stream(self.device)
        self._views = {}                             # start_device: tensor views of the library's buffers
        self.all_slots = None                        # start_device: every rank's slot, on every rank
        self._packed = True                          # (the per-step path of start() packs every step)
        self.mode = "slots"

    def _start_slots(self, local, n_bytes, n_lines):
        """No host round trip and no separate exchange for the sizes: every shard travels as one
        fixed-size transfer into its fixed slot on rank 0 with its {bytes, lines} in front,
        and a kernel of the library (kgm_concat_shards) packs the slots rank-major."""
        slot = (self.k - 1) & 1
        if n_bytes > self.stride:
            raise ValueError("a shard exceeds the exchange buffer")
        done = self.copied[slot]
        # 1. the copy out of the library's buffer, on a stream of its own: the host waits for
        #    this and nothing else.  The slot starts with its own {bytes, lines}.
        with torch.cuda.stream(self.copy_stream):
            for w in self.pending[slot]:
                w.wait()                             # stream-level: the buffers of step k-2 are free
            self.pending[slot] = []
            dst = self.own[slot] if self.rank == 0 else self.xbuf[slot]
            hdr = self.mine_host[slot]
            hdr[0] = n_bytes
            hdr[1] = n_lines
            dst[:16].view(torch.int64).copy_(hdr, non_blocking=True)
            dst[16:16 + n_bytes].copy_(local[:n_bytes], non_blocking=True)
            done.record(self.copy_stream)
        # 2. the shards, on the communication stream, behind the copy
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(done)
            if self.rank == 0:
                self.slots[:16 + n_bytes].copy_(self.own[slot][:16 + n_bytes], non_blocking=True)
                ops = [dist.P2POp(dist.irecv, self.slots[r * self.stride:(r + 1) * self.stride], r, self.group)
                       for r in range(1, self.world)]
                works = dist.batch_isend_irecv(ops) if ops else []
                for w in works:
                    w.wait()                         # the packing kernel below is ordered after the receives
                rc = self._lib.kgm_concat_shards(self.slots.data_ptr(), self.stride, None, self.world,
                                                 self.final.data_ptr(), self.final.numel(),
                                                 self.totals_dev.data_ptr(), self.stream.cuda_stream)
                if rc != 0:
                    raise RuntimeError("kgm_concat_shards failed: %d" % rc)
                fin = torch.cuda.Event()
                fin.record(self.stream)
                self.pending[slot] = [_EventWork(fin)]
                self._packed = True
            else:
                self.pending[slot] = dist.batch_isend_irecv([dist.P2POp(dist.isend, self.xbuf[slot], 0, self.group)])
        done.synchronize()                           # the library may now reuse its output buffer

    def start_device(self, lines_ptr: int, producer):
        """Slots mode with nothing known on the host: ``lines_ptr`` is a device result of the
        library (kgm_encode_resident_async), whose 16 bytes in front hold {n_bytes, n_lines};
        both are written by work already queued on the CUDA stream ``producer``.  The shards
        travel straight out of the library's buffers as ONE fixed-size NCCL all-gather into
        fixed-stride slots, rank-major, each slot = its {n_bytes, n_lines} header + its lines
        (on NVSwitch an all-gather costs rank 0 what a gather would: its ingress is the bound
        either way; it is one collective call instead of world - 1 point-to-point operations,
        no local copy and no host wait are needed, and every rank ends up holding the whole
        result).  The slots ARE the gathered result; ``pack()`` / ``body()`` close the gaps
        between them with kgm_concat_shards when a consumer wants one contiguous buffer --
        not per step: behind the persistent poll kernel of the next step a packing pass over
        every byte only finds a few free SMs and costs more than the step itself.  At least
        ``stride`` bytes must be readable at ``lines_ptr - 16`` (kgm_output_capacity).
        Returns a CUDA event: once it has fired the buffer may be overwritten and the slots
        of this step are complete."""
        if self.mode != "slots":
            raise RuntimeError("start_device needs the slots mode (CUDA, world > 1)")
        self.k += 1
        src = self._views.get(lines_ptr)
        if src is None:
            src = self._views[lines_ptr] = tensor_of_device_buffer(lines_ptr - 16, self.stride, self.device)
        if self.all_slots is None:
            self.all_slots = torch.empty(self.stride * self.world, dtype=torch.uint8, device=self.device)
            if self.rank == 0:
                self.slots = self.all_slots
        ready = torch.cuda.Event()
        ready.record(producer)
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(ready)
            dist.all_gather_into_tensor(self.all_slots, src, group=self.group)
            fin = torch.cuda.Event()
            fin.record(self.stream)
        self._packed = False
        self.pending[(self.k - 1) & 1] = [_EventWork(fin)]
        return fin

    def start(self, local: torch.Tensor, n_bytes: int, n_lines: int):
        self.k += 1
        if self.mode == "slots":
            return self._start_slots(local, n_bytes, n_lines)
        slot = (self.k - 1) & 1
        for w in self.pending[slot]:                 # the exchange buffer of step k-2 is free again
            w.wait()
        self.pending[slot] = []
        counts = gather_counts(n_bytes, n_lines, self.device, self.ctl_group)
        off = slice_offsets(counts)                  # host copy: one sync per step
        sizes = [int(x) for x in off[1:, 0] - off[:-1, 0]]
        if max(sizes) > self.cap:
            raise ValueError("a shard exceeds the exchange buffer")
        xb = self.xbuf[slot]
        mine = int(off[self.rank, 0])
        if self.cuda:
            ev = self.copied[slot]
            with torch.cuda.stream(self.stream):
                if self.rank == 0:
                    self.final[mine:mine + n_bytes].copy_(local[:n_bytes], non_blocking=True)
                else:
                    xb[:n_bytes].copy_(local[:n_bytes], non_blocking=True)
                done = torch.cuda.Event()
                done.record(self.stream)
            if self.mode == "p2p":
                ops = []
                if self.rank == 0:
                    for r in range(1, self.world):
                        if sizes[r]:
                            o = int(off[r, 0])
                            ops.append(dist.P2POp(dist.irecv, self.final[o:o + sizes[r]], r, self.group))
                elif n_bytes:
                    torch.cuda.current_stream(self.device).wait_event(done)
                    ops.append(dist.P2POp(dist.isend, xb[:n_bytes], 0, self.group))
                self.pending[slot] = dist.batch_isend_irecv(ops) if ops else []
            done.synchronize()                       # the library may now reuse its output buffer
        else:
            ops = []
            if self.rank == 0:
                self.final[mine:mine + n_bytes].copy_(local[:n_bytes])
                for r in range(1, self.world):
                    if sizes[r]:
                        o = int(off[r, 0])
                        ops.append(dist.P2POp(dist.irecv, self.final[o:o + sizes[r]], r, self.group))
            elif n_bytes:
                xb[:n_bytes].copy_(local[:n_bytes])
                ops.append(dist.P2POp(dist.isend, xb[:n_bytes], 0, self.group))
            self.pending[slot] = dist.batch_isend_irecv(ops) if ops else []
        self.last_offsets = off
        return off

    def finish(self):
        for slot in (0, 1):
            for w in self.pending[slot]:
                w.wait()
            self.pending[slot] = []
        if self.cuda:
            torch.cuda.synchronize(self.device)
        if self.world > 1:
            dist.barrier(group=self.ctl_group)       # every shard has landed on rank 0

    def pack(self):
        """Rank 0, slots mode: packs the slots of the last step rank-major without gaps into
        ``final`` (kgm_concat_shards, sizes from the slot headers, on the gather's stream)."""
        if self.mode not in ("slots", "push") or self.rank != (self.last_root if self.mode == "push" else 0):
            return
        slots = self.all_slots if self.all_slots is not None else self.slots
        rc = self._lib.kgm_concat_shards(slots.data_ptr(), self.stride, None, self.world,
                                         self.final.data_ptr(), self.final.numel(),
                                         self.totals_dev.data_ptr(), self.stream.cuda_stream)
        if rc != 0:
            raise RuntimeError("kgm_concat_shards failed: %d" % rc)
        self._packed = True

    def body(self):
        """Rank 0: the concatenated bytes of the last finished step."""
        if self.mode in ("slots", "push"):
            if not self._packed:
                self.pack()
                self.stream.synchronize()
            total = int(self.totals_dev[0].item())
            if total < 0:                            # {UINT64_MAX, UINT64_MAX}: a shard did not fit its slot or `final`
                raise RuntimeError("kgm_concat_shards left a shard out (size header beyond the slot or the buffer)")
            return self.final[:total]
        total = int(self.last_offsets[-1, 0])
        return self.final[:total]

    def shard_offsets(self):
        """Byte offset of every rank's shard in the concatenation of the last finished step."""
        if self.mode in ("slots", "push"):
            c = self.slots.view(torch.int64).view(self.world, self.stride // 8)[:, 0].cpu().tolist()
            out = [0]
            for x in c:
                out.append(out[-1] + int(x))
            return out
        return [int(x) for x in self.last_offsets[:, 0].tolist()]


class _EventWork:
    """A CUDA event behind the `wait()` of a distributed work handle (stream-level wait)."""

    def __init__(self, event):
        self.event = event

    def wait(self):
        torch.cuda.current_stream().wait_event(self.event)


class DevicePointer:
    """Zero-copy torch view of a raw device buffer owned by libkgm (CUDA array interface)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (max(int(nbytes), 1),), "typestr": "|u1",
                                         "data": (int(ptr), False), "version": 2}


def tensor_of_device_buffer(ptr: int, nbytes: int, device) -> torch.Tensor:
    return torch.as_tensor(DevicePointer(ptr, nbytes), device=device)
