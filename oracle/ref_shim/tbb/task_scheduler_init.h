// TEST INFRASTRUCTURE ONLY (oracle/): the reference includes this TBB header
// (feature_tracker/optimizer.h:12) but uses nothing from it.
#pragma once
