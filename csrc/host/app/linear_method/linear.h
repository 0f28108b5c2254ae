#pragma once
#include "proto/messages.h"
#include "system/runtime.h"

namespace PS {
namespace LM {
App* createApp(const string& name, const ConfigView& conf);
App* createAppFromText(const string& name, const string& conf_text);
}  // namespace LM
}  // namespace PS
