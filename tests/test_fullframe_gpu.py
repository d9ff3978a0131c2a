"""hotpath.FullFrame: one steady-state frame of DPVO.__call__ (dpvo.py:377-473) from the uint8 image to the updated
window — encoders, gathers, slot writes, edge appends, the update iteration, point cloud, keyframe edits — as stream
launches and as ONE CUDA graph.  The parity of every stage is pinned by its own test file; here: the composition runs,
the graph replays what the stream launches computed, the slot writes land where dpvo.py:435-438 puts them, and the
device edge list stays in its steady state (same count, bit-identical ii/jj/kk) from frame to frame."""
import pytest
import torch

from conftest import load_pkg

pytestmark = pytest.mark.gpu


def test_full_frame_stream_and_graph():
    hotpath = load_pkg("hotpath")
    ff = hotpath.FullFrame(torch.device("cuda", 0), n_frames=12, seed=1, inactive_capacity=400_000)
    assert ff.steady_edges > 0
    e = ff.edges
    before = [t[:ff.steady_edges].clone() for t in (e.active.ii_, e.active.jj_, e.active.kk_)]
    delta, weight = ff.step_frame()
    torch.cuda.synchronize()
    assert torch.isfinite(delta.float()).all() and torch.isfinite(weight.float()).all()
    assert torch.isfinite(ff.poses).all() and torch.isfinite(ff._points).all() and torch.isfinite(ff.motion)
    # image normalisation and slot writes (dpvo.py:389,435-438)
    assert torch.equal(ff.image, 2.0 * (ff.image_u8.float() / 255.0) - 0.5)
    slot, M = (ff.n - 1) % ff.mem, ff.M
    k = ff.n - ff.KEYFRAME_INDEX
    # the keyframe shift moved slot n-1 to n-2 (dpvo.py:287-297) after the write into n-1
    g, i = ff._patches_new
    assert torch.equal(ff.gmap[0, (slot - 1) * M:slot * M], g[0, :, :, :3, :3])
    assert torch.equal(ff.imap[(slot - 1) * M:slot * M], i[0, :, :, 0, 0])
    assert torch.equal(ff.fmap1[0, slot - 1].permute(1, 2, 0), ff._f.nhwc[0])
    assert k < slot
    # motion model: constant-velocity damping keeps a unit quaternion
    assert abs(float(ff._pose_slot[0, 3:].norm()) - 1.0) < 1e-5
    assert float(ff._patch_slot[0, :, 2].min()) == float(ff._patch_slot[0, :, 2].max()) > 0
    # edge list: steady state, bit-identical from frame to frame
    assert e.active.count() == ff.steady_edges
    for a, b in zip(before, (e.active.ii_, e.active.jj_, e.active.kk_)):
        assert torch.equal(a, b[:ff.steady_edges])
    # the graph replays the same frame
    d0, w0 = delta.clone(), weight.clone()
    ff.capture_frame()
    torch.cuda.empty_cache()        # a graph must own or keep alive everything it reads (every later capture does this too)
    d1, w1 = ff.replay_frame()
    torch.cuda.synchronize()
    assert torch.isfinite(d1.float()).all() and d1.shape == d0.shape and w1.shape == w0.shape
    for _ in range(3):
        ff.replay_frame()
    torch.cuda.synchronize()
    assert e.active.count() == ff.steady_edges
    e.check_info()
    # host-buffer variant: uint8 image in, window poses out
    himg, hposes = ff.make_host_image()
    torch.cuda.empty_cache()
    for _ in range(4):
        ff.step_frame_host(himg, hposes)
    torch.cuda.synchronize()
    assert torch.isfinite(hposes).all()
    assert e.active.count() == ff.steady_edges
    e.check_info()


def test_full_frame_pipelined_encoders():
    """make_pipeline / step_pipelined: frame t+1's encoders under frame t's update.  Same image every frame, so the
    ring slot written by each pipelined frame must equal what the in-order frame writes, whatever the overlap."""
    hotpath = load_pkg("hotpath")
    ff = hotpath.FullFrame(torch.device("cuda", 0), n_frames=12, seed=2, inactive_capacity=400_000)
    ff.step_frame()
    torch.cuda.synchronize()
    slot = (ff.n - 1) % ff.mem
    want = ff.fmap1[0, slot - 1].clone()                    # after the keyframe shift the new frame sits in slot n-2
    himg, hposes = ff.make_pipeline()
    torch.cuda.empty_cache()
    for _ in range(7):
        ff.step_pipelined(himg, hposes)
    torch.cuda.synchronize()
    assert torch.equal(ff.fmap1[0, slot - 1], want)
    assert torch.isfinite(hposes).all()
    assert ff.edges.active.count() == ff.steady_edges
    ff.edges.check_info()
