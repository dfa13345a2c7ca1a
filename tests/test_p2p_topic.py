"""The INPUT_DATA topic protocol of the two-process transport (psb200.parallel.p2p), on host shared memory:
a producer process and a consumer process, four slots, back-pressure, in-order delivery, EOF / drained."""
import multiprocessing as mp

import numpy as np

from psb200.parallel import p2p

F, ROWS, SETS, BATCHES = 64, 32, 4, 11


def _batch(k):
    rng = np.random.default_rng(k)
    n = int(rng.integers(1, ROWS + 1))
    nb = int(rng.integers(8, 200))
    return dict(n=n, workers=rng.integers(0, 4, n).astype(np.int32), times=rng.random(n) * 1e3,
                labels=rng.integers(1, 6, n).astype(np.int32),
                bitmap=rng.integers(0, 2 ** 31, (n, F // 32)).astype(np.int32),
                row_off=np.arange(n + 1, dtype=np.int32) * 3, payload=rng.integers(0, 255, nb).astype(np.uint8), nb=nb)


def _consumer(name, nbytes, out):
    lay = p2p.InputTopicLayout(F, ROWS, SETS)
    region = p2p.HostRegion(nbytes, name=name, create=False)
    rd = p2p.InputTopicReader(region, lay)
    got = 0
    try:
        while not rd.drained:
            b = rd.poll()
            if b is None:
                continue
            ref = _batch(b.seq)
            base = lay.set_base(b.seq)
            ok = (b.seq == got and b.n == ref["n"] and b.payload_bytes == ref["nb"]
                  and np.array_equal(b.workers, ref["workers"]) and np.allclose(b.times, ref["times"])
                  and np.array_equal(region.read(base + lay.off_labels, 4 * b.n).view(np.int32), ref["labels"])
                  and np.array_equal(region.read(base + lay.off_bitmap, 4 * b.n * (F // 32)).view(np.int32),
                                     ref["bitmap"].reshape(-1))
                  and np.array_equal(region.read(base + lay.off_payload, b.payload_bytes), ref["payload"])
                  and b.payload == region.address(base + lay.off_payload))
            if not ok:
                out.put(("bad", b.seq))
                return
            got += 1
            rd.mark_consumed(b.seq)
        rd.set_flag(p2p.CTL_WORKER_DONE, 1)
        out.put(("ok", got))
    finally:
        region.close()


def test_layout_slots_do_not_overlap_and_are_aligned():
    lay = p2p.InputTopicLayout(1024, 512, 4)
    offs = [lay.off_hdr, lay.off_workers, lay.off_times, lay.off_labels, lay.off_bitmap, lay.off_row_off, lay.off_payload]
    assert offs == sorted(offs) and all(o % 256 == 0 for o in offs)
    assert lay.off_payload + lay.payload_capacity <= lay.set_bytes
    assert lay.set_base(5) == lay.set_base(1) and lay.set_base(0) >= p2p.CTL_WORDS * 4
    assert lay.bytes == lay.off_sets + 4 * lay.set_bytes


def test_producer_and_consumer_processes_exchange_batches_in_order_with_back_pressure():
    lay = p2p.InputTopicLayout(F, ROWS, SETS)
    region = p2p.HostRegion(lay.bytes)
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    child = ctx.Process(target=_consumer, args=(region.name, lay.bytes, out))
    try:
        wr = p2p.InputTopicWriter(region, lay)
        for k in range(SETS):                                  # fill every slot before the consumer exists
            b = _batch(k)
            assert wr.publish(b["n"], b["workers"], b["times"], b["labels"], b["bitmap"], b["row_off"], b["payload"], b["nb"])
        b = _batch(SETS)
        assert not wr.publish(b["n"], b["workers"], b["times"], b["labels"], b["bitmap"], b["row_off"], b["payload"], b["nb"])
        child.start()
        k = SETS
        while k < BATCHES:
            b = _batch(k)
            if wr.publish(b["n"], b["workers"], b["times"], b["labels"], b["bitmap"], b["row_off"], b["payload"], b["nb"]):
                k += 1
        wr.set_flag(p2p.CTL_EOF)
        assert out.get(timeout=60) == ("ok", BATCHES)
        child.join(timeout=30)
        consumed, done, _ = wr.worker_state()
        assert consumed == BATCHES and done == 1
    finally:
        if child.is_alive():
            child.terminate()
        region.close()
