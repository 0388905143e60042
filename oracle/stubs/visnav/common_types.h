// Stand-in for the reference's include/visnav/common_types.h when its camera_models.h is compiled
// in place for the oracle (TEST INFRASTRUCTURE ONLY).  The real header drags in TBB, OpenCV and
// DBoW2, none of which camera_models.h uses; it only needs <iostream>/<string> to be visible.
#pragma once
#include <iostream>
#include <string>
