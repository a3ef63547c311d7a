// TEST INFRASTRUCTURE - see ../opencv.hpp
#pragma once
#include "../opencv.hpp"
