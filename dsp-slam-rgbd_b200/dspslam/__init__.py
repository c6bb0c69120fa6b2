"""dspslam: Python host side of libdspslam_sm100 (B200-native DSP-SLAM object reconstruction).

`dspslam.engine.Engine` is the batched device engine; `reconstruct.optimizer` (next to this
package) mirrors the reference's `Optimizer` / `MeshExtractor` API on top of it.
"""
