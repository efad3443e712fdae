#pragma once
#include "../mrpt_min.h"
