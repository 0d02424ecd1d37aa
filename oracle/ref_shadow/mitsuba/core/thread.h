#pragma once
#include <mitsuba/core/fresolver.h>
