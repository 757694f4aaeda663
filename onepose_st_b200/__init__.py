"""onepose_st_b200 -- B200-native (sm_100a) implementation of the OnePose++ 2D-3D / LoFTR 2D-2D
matching hot path behind the reference's operator API.  See DESIGN.md."""
__version__ = "0.1.0"
