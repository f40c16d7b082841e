"""The five ORBmatcher::SearchByProjection overloads, both SearchByBoW overloads and DescriptorDistance
of the PRODUCT header (include/ORBmatcher.h, descriptor work on the GPU through liborbfront.so)
against the REFERENCE's own src/ORBmatcher.cc compiled unmodified (oracle/_ref/libmatchref.so), on
identical scenes (tests/projection_cases.py).  Every output (match count, mvpMapPoints, search
radii, matched flags) must be equal."""
import pytest

import projection_cases as pc
import scene_lib as sl

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sides():
    ref = sl.ref_side()
    if ref is None:
        pytest.skip("oracle/_ref/libmatchref.so not built (needs /root/reference at build time)")
    return ref, sl.dropin_side()


LOCAL = [(s, th) for s in (False, True) for th in (1.0, 3.0, 5.0)]
MOTION = [(m, s, th, mo) for (m, s) in ((True, False), (False, True)) for th in (7.0, 15.0, 30.0)
          for mo in ("still", "forward", "backward")]


@pytest.mark.parametrize("stereo,th", LOCAL)
def test_search_local_points(sides, stereo, th):
    pc.search_local_points(*sides, stereo, th)


@pytest.mark.parametrize("mono,stereo,th,motion", MOTION)
def test_track_with_motion_model(sides, mono, stereo, th, motion):
    pc.track_with_motion_model(*sides, mono, stereo, th, motion)


@pytest.mark.parametrize("th,orb_dist", [(10.0, 100), (3.0, 64)])
def test_relocalization_refinement(sides, th, orb_dist):
    pc.relocalization_refinement(*sides, th, orb_dist)


@pytest.mark.parametrize("scale", [1.0, 1.25])
def test_loop_closure_sim3(sides, scale):
    pc.loop_closure_sim3(*sides, scale)


@pytest.mark.parametrize("th", [1.0, 15.0])
def test_map_wide_pose_candidate(sides, th):
    pc.map_wide_pose_candidate(*sides, th)


def test_bow_and_distance_against_the_reference_source(sides):
    pc.bow_and_distance(*sides)


@pytest.mark.parametrize("stereo,th", [(False, 3.0), (True, 3.0), (True, 6.0)])
def test_fuse_into_neighbor_keyframe(sides, stereo, th):
    pc.fuse_neighbors(*sides, stereo, th)


@pytest.mark.parametrize("scale", [1.0, 1.2])
def test_fuse_loop_points(sides, scale):
    pc.fuse_loop(*sides, scale)


@pytest.mark.parametrize("s12", [1.0, 1.15])
def test_search_by_sim3(sides, s12):
    pc.sim3_mutual(*sides, s12)


@pytest.mark.parametrize("stereo,only_stereo,check_ori", [(False, False, True), (True, False, True), (True, True, True),
                                                          (False, False, False)])
def test_search_for_triangulation(sides, stereo, only_stereo, check_ori):
    pc.triangulation(*sides, stereo, only_stereo, check_ori)


@pytest.mark.parametrize("window,ratio,check_ori", [(100, 0.9, True), (100, 0.9, False), (30, 0.7, True)])
def test_search_for_initialization(sides, window, ratio, check_ori):
    pc.initialization(*sides, window, ratio, check_ori)
