"""Enumerations of the reference's public surface (gym_pybullet_drones/utils/enums.py:3-49), same names and
values, so that constructor calls written for the reference keep working."""
from enum import Enum


class DroneModel(Enum):
    CF2X = "cf2x"
    CF2P = "cf2p"
    RACE = "racer"


class Physics(Enum):
    PYB = "pyb"
    DYN = "dyn"
    PYB_GND = "pyb_gnd"
    PYB_DRAG = "pyb_drag"
    PYB_DW = "pyb_dw"
    PYB_GND_DRAG_DW = "pyb_gnd_drag_dw"


class ImageType(Enum):
    RGB = 0
    DEP = 1
    SEG = 2
    BW = 3


class ActionType(Enum):
    RPM = "rpm"
    PID = "pid"
    VEL = "vel"
    ONE_D_RPM = "one_d_rpm"
    ONE_D_PID = "one_d_pid"


class ObservationType(Enum):
    KIN = "kin"
    RGB = "rgb"
    COKIN = "cokin"


# ChsAct values of include/chs.h
ACT_CODE = {ActionType.VEL: 0, ActionType.RPM: 1, ActionType.PID: 2, ActionType.ONE_D_RPM: 3,
            ActionType.ONE_D_PID: 4, "vel": 0, "rpm": 1, "pid": 2, "one_d_rpm": 3, "one_d_pid": 4}
ACT_WIDTH = {0: 4, 1: 4, 2: 3, 3: 1, 4: 1}
