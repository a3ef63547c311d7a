// TEST INFRASTRUCTURE - see ceres.h in this directory
#pragma once
#include "ceres.h"
