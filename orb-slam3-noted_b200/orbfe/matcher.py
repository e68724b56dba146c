"""ORBmatcher — Python mirror of ORB_SLAM3::ORBmatcher (include/ORBmatcher.h:35-111) over the C ABI."""


class FeatureVectorCSR:
    pass


class ORBmatcher:
    TH_LOW = 50
    TH_HIGH = 100
    HISTO_LENGTH = 30
