"""The known answers of tests/test_oracle_fixed_function.py — coverage sets from the Vulkan rules restated in exact rational
arithmetic, top-left ties, slivers, winding and culling, LESS_OR_EQUAL in submission order, depth planes, near / far clipping,
the scissor, attachment conversions, the additive blend, the Transparent pass's depth state, the shadow pass's cull mode — asked
of the KERNELS directly: their host emulation on the CPU, libvkr on the B200.  The parity tests hold the kernels to the oracle and
that file holds the oracle to the rules; this one closes the triangle without going through the oracle at all.
The texture-unit cases (mip blit, REPEAT bilinear, NEAREST, trilinear level selection, cube faces, seamless cube edges) go
through probe entry points that only the emulation has beside the oracle: they run on the emulated kernels."""
import pytest

import test_oracle_fixed_function as ff

CASES = [
    ff.test_rectangle_on_pixel_corners_covers_exactly_its_pixels,
    ff.test_rectangle_on_pixel_centres_keeps_top_and_left_drops_bottom_and_right,
    ff.test_shared_diagonal_through_pixel_centres_is_covered_exactly_once,
    ff.test_fan_vertex_on_a_pixel_centre_is_covered_exactly_once,
    ff.test_slivers_and_degenerates,
    ff.test_back_faces_are_culled_and_front_is_counter_clockwise,
    ff.test_depth_test_is_less_or_equal_in_submission_order,
    ff.test_depth_zero_and_one_are_kept,
    ff.test_depth_is_the_plane_through_the_vertices_at_the_sample,
    ff.test_random_triangles_match_the_rules,
    ff.test_near_plane_clip_of_one_triangle,
    ff.test_far_plane_clip,
    ff.test_scissor_keeps_exactly_its_pixels,
    ff.test_colour_attachment_conversions,
    ff.test_additive_blend_sums_in_the_f16_attachment,
    ff.test_transparent_pass_tests_depth_but_does_not_write_it,
    ff.test_shadow_pass_culls_front_faces_and_stores_the_plane_depth,
]


TEXTURE_CASES = [
    ff.test_mip_blit_is_a_linear_light_box_filter,
    ff.test_bilinear_repeat_sample_across_the_wrap,
    ff.test_trilinear_level_selection,
    ff.test_cube_face_selection_and_orientation,
    ff.test_cube_filtering_is_seamless_across_all_twelve_edges,
]


@pytest.mark.parametrize("case", CASES + TEXTURE_CASES, ids=lambda f: f.__name__[5:])
def test_rule_on_the_emulated_kernels(case, emu_lib, monkeypatch):
    monkeypatch.setattr(ff, "BACKEND_PREFIX", "emu")
    case(emu_lib)


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES, ids=lambda f: f.__name__[5:])
def test_rule_on_the_device(case, vkr_lib, monkeypatch):
    monkeypatch.setattr(ff, "BACKEND_PREFIX", "vkr")
    case(vkr_lib)
