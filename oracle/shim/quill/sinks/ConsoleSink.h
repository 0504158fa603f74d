#pragma once
#include "quill/Logger.h"
