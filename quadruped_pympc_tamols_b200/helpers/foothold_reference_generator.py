"""Batched drop-in for quadruped_pympc/helpers/foothold_reference_generator.py (class FootholdReferenceGenerator,
compute_footholds_reference :53-157) -- SURVEY.md section 8(f) row N3: the seed footholds of R robots in one kernel
launch, and (optionally, same launch) the tamols_leg_job records that cut every leg's candidate window out of ONE
global height map, so that states -> seeds -> jobs -> TAMOLS scores -> footholds never leaves the device.

The moving-average velocity deque (:95-96) is a device ring per robot, advanced by every call like upstream.  float64.
No CPU path: inputs and outputs are torch CUDA tensors, the work is the kernel behind tamols_seed_jobs (include/tamols.h).
"""
import numpy as np

from .. import _tamols_cabi as tc


class FootholdReferenceGenerator:
    """R foothold generators on the GPU of `handle` (a quadruped_pympc_tamols_b200._tamols_cabi.Handle)."""

    def __init__(self, stance_time, lift_off_positions, vel_moving_average_length=20, hip_height=None, num_robots=1,
                 handle=None, gravity_constant=9.81):
        import torch  # device buffers only
        if handle is None:
            raise tc.TamolsError("FootholdReferenceGenerator needs the tamols handle whose stream it runs on "
                                 "(there is no CPU path)")
        if hip_height is None:
            raise ValueError("hip_height is required (upstream multiplies None by 1.5 and fails later)")
        self._torch, self._h = torch, handle
        self._dev = torch.device("cuda", int(handle.device))
        self.num_robots = R = int(num_robots)
        self.hist_len = int(vel_moving_average_length)
        self.hip_height = float(hip_height)
        self.hip_offset = 0.1  # :44
        self.gravity_constant = float(gravity_constant)
        f64 = torch.float64
        self.stance_time = self._up(np.broadcast_to(np.asarray(stance_time, np.float64), (R,)))
        lo = np.asarray(lift_off_positions, np.float64).reshape(-1, 4, 3)
        self.lift_off_positions = self._up(np.broadcast_to(lo, (R, 4, 3)))  # [R,4,3]; only z enters the seeds (:150)
        self.com_pos_offset_b = torch.zeros((R, 3), dtype=f64, device=self._dev)
        self._hist = torch.zeros((R, self.hist_len, 2), dtype=f64, device=self._dev)
        self._count = torch.zeros((R,), dtype=torch.int32, device=self._dev)
        self.last_reference_footholds = torch.zeros((R, 4, 3), dtype=f64, device=self._dev)
        self._lift_z = None

    def _up(self, a):
        return self._torch.as_tensor(np.array(a, np.float64), dtype=self._torch.float64, device=self._dev)

    def _t(self, a, shape):
        t = a if self._torch.is_tensor(a) else self._up(a)
        t = t.to(self._dev, self._torch.float64).reshape(shape).contiguous()
        return t

    def compute_footholds_reference(self, base_position, base_ori_euler_xyz, base_lin_vel, ref_base_lin_vel,
                                    hips_position, com_height_nominal, feet_position=None, current_contact=None,
                                    global_map=None, window=None, jobs_out=None):
        """[R,4,3] float64 CUDA tensor of reference footholds (:84-157).  base_lin_vel / ref_base_lin_vel are [R,3]
        world-frame velocities (x, y enter the seeds like upstream's base_xy_lin_vel arguments).

        With `global_map` = (rows, cols, origin_xy, row_axis_xy, col_axis_xy), `window` = (rows, cols) and `jobs_out` =
        a uint8 CUDA tensor of 4 R * 136 bytes, the same launch also writes the tamols_leg_job records (feet_position
        [R,4,3] and current_contact [R,4] give the stability inputs, visual_foothold_adaptation.py:631-664)."""
        t, R = self._torch, self.num_robots
        keep = [self._t(base_position, (R, 3)), self._t(base_ori_euler_xyz, (R, 3)), self._t(base_lin_vel, (R, 3)),
                self._t(ref_base_lin_vel, (R, 3)), self._t(hips_position, (R, 4, 3)),
                self._t(np.broadcast_to(np.asarray(com_height_nominal, np.float64), (R,))
                        if not t.is_tensor(com_height_nominal) else com_height_nominal, (R,))]
        self._lift_z = self.lift_off_positions[:, :, 2].contiguous()
        b = tc.TamolsSeedBatch(num_robots=R, hist_len=self.hist_len, d_base_position=keep[0].data_ptr(),
                               d_base_euler=keep[1].data_ptr(), d_base_lin_vel=keep[2].data_ptr(),
                               d_ref_lin_vel=keep[3].data_ptr(), d_hips=keep[4].data_ptr(),
                               d_com_height=keep[5].data_ptr(), d_stance_time=self.stance_time.data_ptr(),
                               d_lift_off_z=self._lift_z.data_ptr(), d_com_offset_b=self.com_pos_offset_b.data_ptr(),
                               d_vel_hist=self._hist.data_ptr(), d_hist_count=self._count.data_ptr(),
                               hip_height=self.hip_height, hip_offset=self.hip_offset, gravity=self.gravity_constant,
                               d_ref_feet=self.last_reference_footholds.data_ptr())
        if jobs_out is not None:
            rows, cols, origin, row_axis, col_axis = global_map
            assert jobs_out.numel() * jobs_out.element_size() >= 4 * R * tc.JOB_DTYPE.itemsize
            b.d_jobs = jobs_out.data_ptr()
            b.map_rows, b.map_cols, b.win_rows, b.win_cols = int(rows), int(cols), int(window[0]), int(window[1])
            for k in range(2):
                b.map_origin[k], b.map_row_axis[k], b.map_col_axis[k] = float(origin[k]), float(row_axis[k]), float(col_axis[k])
            if feet_position is not None:
                keep.append(self._t(feet_position, (R, 4, 3)))
                b.d_feet = keep[-1].data_ptr()
            if current_contact is not None:
                c = current_contact if t.is_tensor(current_contact) else t.as_tensor(np.array(current_contact, np.uint8))
                keep.append(c.to(self._dev, t.uint8).reshape(R, 4).contiguous())
                b.d_contact = keep[-1].data_ptr()
        t.cuda.current_stream(self._dev).synchronize()  # inputs were produced on torch's stream
        self._h.seed_jobs(b)
        self._keep = keep  # the launch is asynchronous: keep its inputs alive until the next call
        return self.last_reference_footholds
