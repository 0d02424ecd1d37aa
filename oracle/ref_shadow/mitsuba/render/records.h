#pragma once
#include <mitsuba/render/interaction.h>
