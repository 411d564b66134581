"""EChannel / ChannelMask constants (inc/ChannelMap.h:38-147). Pure Python: importable without the native library."""


class EChannel:
    """inc/ChannelMap.h:38-79"""
    FrontLeft, FrontRight, FrontCenter, LFE = 1 << 0, 1 << 1, 1 << 2, 1 << 3
    SideLeft, SideRight, FrontLeftCenter, FrontRightCenter = 1 << 4, 1 << 5, 1 << 6, 1 << 7
    BackLeft, BackRight, BackCenter, WideLeft, WideRight = 1 << 8, 1 << 9, 1 << 10, 1 << 11, 1 << 12
    TopCenter, TopFrontLeft, TopFrontCenter, TopFrontRight = 1 << 13, 1 << 14, 1 << 15, 1 << 16
    TopSideLeft, TopSideRight, TopBackLeft, TopBackCenter, TopBackRight = 1 << 17, 1 << 18, 1 << 19, 1 << 20, 1 << 21


class ChannelMask:
    """inc/ChannelMap.h:87-147"""
    E = EChannel
    Invalid = 0
    Mono = E.FrontCenter
    Stereo = E.FrontLeft | E.FrontRight
    LCR = Stereo | E.FrontCenter
    LRS = Stereo | E.BackCenter
    LCRS = LCR | E.BackCenter
    Quad = Stereo | E.BackLeft | E.BackRight
    Octagonal = LCR | E.BackLeft | E.BackRight | E.BackCenter | E.WideLeft | E.WideRight
    Surround_4_1 = Quad | E.LFE
    Surround_5_0 = LCR | E.BackLeft | E.BackRight
    Surround_5_1 = Surround_5_0 | E.LFE
    Surround_6_0 = Surround_5_0 | E.BackCenter
    Surround_6_1 = Surround_6_0 | E.LFE
    Surround_7_0 = Surround_5_0 | E.SideLeft | E.SideRight
    Surround_7_1 = Surround_7_0 | E.LFE
    Surround_5_0_2 = Surround_5_0 | E.TopSideLeft | E.TopSideRight
    Surround_5_1_2 = Surround_5_0_2 | E.LFE
    Surround_5_0_4 = Surround_5_0 | E.TopFrontLeft | E.TopFrontRight | E.TopBackLeft | E.TopBackRight
    Surround_5_1_4 = Surround_5_0_4 | E.LFE
    Surround_7_0_2 = Surround_7_0 | E.TopSideLeft | E.TopSideRight
    Surround_7_1_2 = Surround_7_0_2 | E.LFE
    Surround_7_0_4 = Surround_7_0 | E.TopFrontLeft | E.TopFrontRight | E.TopBackLeft | E.TopBackRight
    Surround_7_1_4 = Surround_7_0_4 | E.LFE
    Surround_7_0_6 = Surround_7_0_4 | E.TopSideLeft | E.TopSideRight
    Surround_7_1_6 = Surround_7_0_6 | E.LFE
    Surround_9_0_4 = Surround_7_0_4 | E.WideLeft | E.WideRight
    Surround_9_1_4 = Surround_9_0_4 | E.LFE
    Surround_9_0_6 = Surround_9_0_4 | E.TopSideLeft | E.TopSideRight
    Surround_9_1_6 = Surround_9_0_6 | E.LFE
