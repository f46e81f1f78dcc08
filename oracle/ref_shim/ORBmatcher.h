// Include shim: lets the reference's own ORBmatcher.h be used verbatim while keeping its heavy dependencies out.
// The reference's ORBmatcher.cc does `#include "ORBmatcher.h"`; with -I oracle/ref_shim first on the include path it
// lands here.  We pre-define the include guards of the reference's MapPoint.h / KeyFrame.h / Frame.h (which pull in
// libtorch, ROS and g2o -- none installed) so that they expand to nothing, provide slam_shim.h instead, and then
// include the reference header from where it lies.  TEST INFRASTRUCTURE ONLY.
#pragma once
#define MAPPOINT_H
#define KEYFRAME_H
#define FRAME_H
#include "slam_shim.h"
#ifndef ASD_REF_ORBMATCHER_H
#error "pass -DASD_REF_ORBMATCHER_H='\"/root/reference/src/vslam/include/ORBmatcher.h\"'"
#endif
#include ASD_REF_ORBMATCHER_H
