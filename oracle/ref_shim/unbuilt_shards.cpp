// Build shim (oracle/Makefile.refcuda): the reference's kernel-template shards that this build does
// NOT compile (minutes of nvcc each) are defined here as functions that throw, so that the library
// links and a shape that dispatches to one of them fails loudly.  -DHAVE_<shard file name> marks the
// shards that were compiled from the reference's own .cu files.
#include <string>

#include "dau_conv/dau_conv_impl/dau_conv_backward.hpp"
#include "dau_conv/dau_conv_impl/dau_conv_forward.hpp"
#include "dau_conv/util/common.hpp"

namespace DAUConvNet {

#define FWD_STUB(fn) \
    void fn(int, int, int, int, int, int, int, DAUConvForward<float>::CUDAParams&) { \
        throw DAUException(std::string("reference shard not built into oracle/_ref: ") + #fn); }
#define BWD_STUB(fn) \
    void fn(int, int, int, bool, int, bool, bool, DAUConvBackward<float>::CUDAParams&) { \
        throw DAUException(std::string("reference shard not built into oracle/_ref: ") + #fn); }

#ifndef HAVE_dau_conv_forward_off4_s0_f0
FWD_STUB(DAUConv_forward_float_off_4_single_feat_0_single_subfeat_0)
#endif
#ifndef HAVE_dau_conv_forward_off4_s0_f1
FWD_STUB(DAUConv_forward_float_off_4_single_feat_0_single_subfeat_1)
#endif
#ifndef HAVE_dau_conv_forward_off4_s1_f0
FWD_STUB(DAUConv_forward_float_off_4_single_feat_1_single_subfeat_0)
#endif
#ifndef HAVE_dau_conv_forward_off4_s1_f1
FWD_STUB(DAUConv_forward_float_off_4_single_feat_1_single_subfeat_1)
#endif
#ifndef HAVE_dau_conv_forward_off8_s0_f0
FWD_STUB(DAUConv_forward_float_off_8_single_feat_0_single_subfeat_0)
#endif
#ifndef HAVE_dau_conv_forward_off8_s0_f1
FWD_STUB(DAUConv_forward_float_off_8_single_feat_0_single_subfeat_1)
#endif
#ifndef HAVE_dau_conv_forward_off8_s1_f0
FWD_STUB(DAUConv_forward_float_off_8_single_feat_1_single_subfeat_0)
#endif
#ifndef HAVE_dau_conv_forward_off8_s1_f1
FWD_STUB(DAUConv_forward_float_off_8_single_feat_1_single_subfeat_1)
#endif
#ifndef HAVE_dau_conv_forward_off16_s0_f1
FWD_STUB(DAUConv_forward_float_off_16_single_feat_0_single_subfeat_1)
#endif
#ifndef HAVE_dau_conv_forward_off16_s1_f1
FWD_STUB(DAUConv_forward_float_off_16_single_feat_1_single_subfeat_1)
#endif
#ifndef HAVE_dau_conv_forward_off32_s1_f1
FWD_STUB(DAUConv_forward_float_off_32_single_feat_1_single_subfeat_1)
#endif

#ifndef HAVE_dau_conv_backward_patch_1x1
BWD_STUB(DAUConv_backward_multi_subfeatures_patch_1x1)
#endif
#ifndef HAVE_dau_conv_backward_patch_8x8
BWD_STUB(DAUConv_backward_multi_subfeatures_patch_8x8)
#endif
#ifndef HAVE_dau_conv_backward_patch_8x16
BWD_STUB(DAUConv_backward_multi_subfeatures_patch_8x16)
#endif
#ifndef HAVE_dau_conv_backward_patch_8x32
BWD_STUB(DAUConv_backward_multi_subfeatures_patch_8x32)
#endif
#ifndef HAVE_dau_conv_backward_patch_8x64
BWD_STUB(DAUConv_backward_multi_subfeatures_patch_8x64)
#endif
#ifndef HAVE_dau_conv_backward_patch_16x8
BWD_STUB(DAUConv_backward_multi_subfeatures_patch_16x8)
#endif
#ifndef HAVE_dau_conv_backward_patch_16x16
BWD_STUB(DAUConv_backward_multi_subfeatures_patch_16x16)
#endif
#ifndef HAVE_dau_conv_backward_patch_16x32
BWD_STUB(DAUConv_backward_multi_subfeatures_patch_16x32)
#endif
#ifndef HAVE_dau_conv_backward_patch_16x64
BWD_STUB(DAUConv_backward_multi_subfeatures_patch_16x64)
#endif
#ifndef HAVE_dau_conv_backward_patch_32x8
BWD_STUB(DAUConv_backward_multi_subfeatures_patch_32x8)
#endif
#ifndef HAVE_dau_conv_backward_patch_32x16
BWD_STUB(DAUConv_backward_multi_subfeatures_patch_32x16)
#endif
#ifndef HAVE_dau_conv_backward_patch_32x32
BWD_STUB(DAUConv_backward_multi_subfeatures_patch_32x32)
#endif
#ifndef HAVE_dau_conv_backward_patch_32x64
BWD_STUB(DAUConv_backward_multi_subfeatures_patch_32x64)
#endif
#ifndef HAVE_dau_conv_backward_patch_64x8
BWD_STUB(DAUConv_backward_multi_subfeatures_patch_64x8)
#endif
#ifndef HAVE_dau_conv_backward_patch_64x16
BWD_STUB(DAUConv_backward_multi_subfeatures_patch_64x16)
#endif
#ifndef HAVE_dau_conv_backward_patch_64x32
BWD_STUB(DAUConv_backward_multi_subfeatures_patch_64x32)
#endif
#ifndef HAVE_dau_conv_backward_patch_64x64
BWD_STUB(DAUConv_backward_multi_subfeatures_patch_64x64)
#endif

}  // namespace DAUConvNet
