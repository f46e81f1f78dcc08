// stub: the reference KeyFrameDatabase.cc includes "KeyFrame.h"; the shim types live in oracle/ref_shim/kfdb/KeyFrameDatabase.h
#pragma once
