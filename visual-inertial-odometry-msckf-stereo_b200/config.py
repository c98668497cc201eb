"""Hot-path configuration: the fields of the reference's ``ConfigEuRoC`` /
``OptimizationConfigEuRoC`` (config.py:5-124) that the MSCKF back end reads, with the
reference's attribute names and default values.  The front-end (cv2) fields are left
out -- the KLT/FAST front end is out of scope -- so this module does not import cv2.

Two values the reference hard-codes are exposed here as overridable attributes with
the reference's literals as defaults: the 1500-row cap of ``remove_lost_features``
(msckf.py:678) and the 0.05 chi-square tail (msckf.py:123).
"""
import numpy as np


class OptimizationConfigEuRoC(object):
    """config.py:5-15"""

    def __init__(self):
        self._vio_translation_threshold__ = -1.0
        self._vio_huber_epsilon__ = 0.01
        self._vio_estimation_precision__ = 5e-7
        self._vio_initial_damping__ = 1e-3
        self._vio_outer_loop_max_iteration__ = 5
        self._vio_inner_loop_max_iteration__ = 5


class ConfigEuRoC(object):
    """config.py:17-124 (back-end fields)"""

    def __init__(self):
        self._vio_optimization_config__ = OptimizationConfigEuRoC()
        self._vio_gravity_acc__ = 9.81
        self._vio_gravity__ = np.array([0.0, 0.0, -self._vio_gravity_acc__])
        self._vio_frame_rate__ = 20
        self._vio_max_cam_state_size__ = 20
        self._vio_position_std_threshold__ = 8.0
        # config.py:67-69 -- unused by the reference (msckf.py:717 hard-codes the same numbers)
        self._vio_rotation_threshold__ = 0.2618
        self._vio_translation_threshold__ = 0.4
        self._vio_tracking_rate_threshold__ = 0.5
        self._vio_gyro_noise__ = 0.005 ** 2
        self._vio_acc_noise__ = 0.05 ** 2
        self._vio_gyro_bias_noise__ = 0.001 ** 2
        self._vio_acc_bias_noise__ = 0.01 ** 2
        self._vio_observation_noise__ = 0.035 ** 2
        self._vio_velocity__ = np.zeros(3)
        self._vio_velocity_cov__ = 0.25
        self._vio_gyro_bias_cov__ = 0.01
        self._vio_acc_bias_cov__ = 0.01
        self._vio_extrinsic_rotation_cov__ = 3.0462e-4
        self._vio_extrinsic_translation_cov__ = 2.5e-5
        self._vio_T_imu_cam0__ = np.array([
            [0.014865542981794, 0.999557249008346, -0.025774436697440, 0.065222909535531],
            [-0.999880929698575, 0.014967213324719, 0.003756188357967, -0.020706385492719],
            [0.004140296794224, 0.025715529947966, 0.999660727177902, -0.008054602460030],
            [0, 0, 0, 1.000000000000000]])
        self._vio_T_cn_cnm1__ = np.array([
            [0.999997256477881, 0.002312067192424, 0.000376008102415, -0.110073808127187],
            [-0.002317135723281, 0.999898048506644, 0.014089835846648, 0.000399121547014],
            [-0.000343393120525, -0.014090668452714, 0.999900662637729, -0.000853702503357],
            [0, 0, 0, 1.000000000000000]])
        self._vio_T_imu_body__ = np.eye(4)
        # camera calibration of the EuRoC sensor (config.py:99-121), used by image.undistort_stereo
        self._vio_cam0_distortion_model__ = 'radtan'
        self._vio_cam0_distortion_coeffs__ = np.array([-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05])
        self._vio_cam0_intrinsics__ = np.array([458.654, 457.296, 367.215, 248.375])
        self._vio_cam0_resolution__ = np.array([752, 480])
        self._vio_cam1_distortion_model__ = 'radtan'
        self._vio_cam1_distortion_coeffs__ = np.array([-0.28368365, 0.07451284, -0.00010473, -3.55590700e-05])
        self._vio_cam1_intrinsics__ = np.array([457.587, 456.134, 379.999, 255.238])
        self._vio_cam1_resolution__ = np.array([752, 480])
        # literals of the reference, overridable here
        self.row_cap = 1500            # msckf.py:678
        self.chi2_tail = 0.05          # msckf.py:123
