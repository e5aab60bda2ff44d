"""Host-side helpers of the synchronous fast path that can be checked without a GPU."""
import torch

from elephas_b200.parallel import sync_fast


class _TwoRanks:
    """Stands in for the runtime: collects the rows 'sent' by two emulated ranks, then answers with all of them."""

    def __init__(self):
        self.rows = []

    def all_gather_rows(self, row):
        self.rows.append(row.clone())
        return torch.stack(self.rows)


def test_histories_travel_as_one_row_per_rank():
    keys = ['loss', 'acc', 'val_loss', 'val_acc', 'batch_loss']
    h0 = {'loss': [0.5, 0.25], 'acc': [0.1, 0.2], 'batch_loss': [1.0, 0.75, 0.5, 0.3]}
    rt = _TwoRanks()
    first = sync_fast._gather_histories(rt, None, keys, 4)          # rank 0 had nothing to train on
    assert first == [None]
    both = sync_fast._gather_histories(rt, h0, keys, 4)
    assert both[0] is None and both[1] == h0 and list(both[1]) == list(h0)
    h2 = {'loss': [1e-30, 3.0], 'acc': [], 'val_loss': [2.0, 1.0]}   # an empty list is not an absent key
    assert sync_fast._gather_histories(_TwoRanks(), h2, keys, 4) == [h2]


def test_nvls_master_gather_only_when_the_buckets_tile_the_buffer():
    """The multimem.st gather of the sharded fp32 master copies the whole staging buffer over the master: only valid when
    every element belongs to exactly one bucket (then the ranks' owned slices cover the buffer)."""
    class _Eng:
        def __init__(self, ranges, numel):
            self._ranges, self.w32 = ranges, torch.zeros(numel)

        def bucket_ranges(self):
            return self._ranges

    class _Rt:
        rank, world_size = 1, 4

    def ctx(ranges, numel):
        c = sync_fast.SyncContext.__new__(sync_fast.SyncContext)
        c.eng, c.rt = _Eng(ranges, numel), _Rt()
        return c

    tiled = ctx([(128, 64), (0, 128), (192, 64)], 256)            # emitted last layer first, vectors last: any order
    assert tiled._ranges_cover_master()
    owned = tiled._owned_ranges()
    assert owned == [(128 + 16, 128 + 32), (32, 64), (192 + 16, 192 + 32)]      # slice 1 of 4 of every bucket
    assert not ctx([(0, 128), (192, 64)], 256)._ranges_cover_master()          # a gap
    assert not ctx([(0, 128), (128, 64)], 256)._ranges_cover_master()          # the tail is nobody's
    assert not ctx([(0, 132), (132, 124)], 256)._ranges_cover_master()         # slices must be groups of 8 elements


def test_packed_histories_round_trip_for_arbitrary_shapes():
    """Property: whatever subset of the keys a rank's history holds and however long its lists are (up to the width), every
    rank gets back exactly what was sent, None included, keys in canonical order."""
    from hypothesis import given, settings
    from hypothesis import strategies as st

    keys = ['loss', 'acc', 'val_loss', 'val_acc', 'batch_loss']
    values = st.floats(allow_nan=False, allow_infinity=False, width=64)
    history = st.one_of(st.none(), st.dictionaries(st.sampled_from(keys), st.lists(values, max_size=6), max_size=5))

    @settings(max_examples=60, deadline=None, derandomize=True)
    @given(st.lists(history, min_size=1, max_size=4))
    def check(per_rank):
        rt = _TwoRanks()
        result = None
        for hist in per_rank:   # rank after rank "sends" its row; the last call sees all of them
            ordered = None if hist is None else {k: hist[k] for k in keys if k in hist}
            result = sync_fast._gather_histories(rt, ordered, keys, 6)
        expected = [None if h is None else {k: h[k] for k in keys if k in h} for h in per_rank]
        assert result == expected

    check()
