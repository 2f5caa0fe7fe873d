"""Shared builders for the matcher / stereo parity tests: the same seeded data fed to the oracle views
(oracle.pyoracle) and to the product views (my_orb_slam2_b200.frontend)."""
import numpy as np

from my_orb_slam2_b200 import synth


def extract_with_oracle(oracle, shape="kitti", seed=0, pair=False, small=None):
    s = synth.SETTINGS[shape]
    w, h = small or synth.SHAPES[shape]
    nf = s["nfeatures"] if small is None else max(100, int(s["nfeatures"] * (w * h) / (synth.SHAPES[shape][0] * synth.SHAPES[shape][1])))
    params = (nf, s["scale_factor"], s["nlevels"], s["ini_th"], s["min_th"])
    if pair:
        left, right = synth.synth_pair(w, h, seed)
        exl, exr = oracle.Extractor(*params), oracle.Extractor(*params)
        return params, (left, right), (exl, exr), (exl.extract(left), exr.extract(right))
    img = synth.synth_image(w, h, seed)
    ex = oracle.Extractor(*params)
    return params, img, ex, ex.extract(img)


def points_view(mod, p):
    f = mod.make_points3d_view if hasattr(mod, "make_frame_view") else mod.points3d_view
    return f(p["valid"], p["world_pos"], p["normal"], p["max_distance"], p["min_distance"], p["desc"], p["angle"])


def projection(mod, pr, ow=None, is_sim3=None):
    f = mod.make_projection if hasattr(mod, "make_frame_view") else mod.projection
    return f(pr["tcw"], pr["fx"], pr["fy"], pr["cx"], pr["cy"], pr["bf"], pr["log_scale_factor"],
             pr["is_sim3"] if is_sim3 is None else is_sim3, ow if ow is not None else pr["ow"])


def sim3_pair(mod, s):
    f = mod.make_sim3_pair if hasattr(mod, "make_frame_view") else mod.sim3_pair
    return f(s["t1w"], s["t2w"], s["s12"], s["r12"], s["t12"], s["fx"], s["fy"], s["cx"], s["cy"], s["log_sf1"],
             s["log_sf2"])


def views(mod, frame=None, mps=None, last=None, cam=None, kf=None):
    """Build views with either oracle.pyoracle (make_*_view) or my_orb_slam2_b200.frontend (*_view)."""
    is_oracle = hasattr(mod, "make_frame_view")
    out = []
    if frame is not None:
        f = mod.make_frame_view if is_oracle else mod.frame_view
        out.append(f(frame["keys_un"], frame["u_right"], frame["desc"], frame["mp_obs"], frame["bounds"],
                     frame["scale_factors"]))
    if mps is not None:
        f = mod.make_mappoints_view if is_oracle else mod.mappoints_view
        out.append(f(mps["proj_x"], mps["proj_y"], mps["proj_xr"], mps["level"], mps["view_cos"], mps["in_view"],
                     mps["obs"], mps["desc"]))
    if last is not None:
        f = mod.make_lastframe_view if is_oracle else mod.lastframe_view
        out.append(f(last["valid"], last["world_pos"], last["obs"], last["desc"], last["octave"], last["angle"],
                     last["tcw"]))
    if cam is not None:
        f = mod.make_camera_pose if is_oracle else mod.camera_pose
        out.append(f(cam["tcw"], cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["mbf"], cam["mb"]))
    if kf is not None:
        f = mod.make_keyframe_view if is_oracle else mod.keyframe_view
        out.append(f(kf["keys_un"], kf["u_right"], kf["desc"], kf["has_mp"], kf["node_ids"], kf["node_offsets"],
                     kf["node_features"], kf["scale_factors"], kf["level_sigma2"]))
    return out if len(out) > 1 else out[0]
