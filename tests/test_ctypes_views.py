"""the zero-copy result path of the ctypes stub (FileResults) against the copying one"""
import pytest


@pytest.mark.gpu
def test_file_results_views_equal_copies(tmpdir_path):
    import os
    import avr_ctypes as A
    from conftest import gen_clip
    blobs = [gen_clip(os.path.join(tmpdir_path, f"c{i}.h264"), seed=70 + i, mb_w=8, mb_h=6, frames=2, slices_per_frame=1 + i) for i in range(3)]
    blobs.append(b"")                                   # empty input -> one empty literal block
    with A.Context(0) as ctx:
        comp, _ = A.compress_files(ctx, blobs)
        with A.compress_files(ctx, blobs, copy=False)[0] as r:
            assert [bytes(v) for v in r.views] == comp
            with A.decompress_files(ctx, r.views, copy=False)[0] as back:      # views of one call feed the next
                assert [bytes(v) for v in back.views] == blobs
