// TEST SHIM -- not the reference (see event/event.h).  Stand-in for src/mef/openpsa/settings.h (this one reference header
// compiles standalone, but it is not copied: only the accessors the adapter calls).
#pragma once
namespace mef::openpsa {

class Settings {
  public:
    int num_trials() const { return num_trials_; }                      // settings.h:118
    Settings& num_trials(int n) { num_trials_ = n; return *this; }      // settings.h:127
    int seed() const { return seed_; }                                  // settings.h:178
    Settings& seed(int s) { seed_ = s; return *this; }                  // settings.h:184
    bool ccf_analysis() const { return ccf_analysis_; }                 // settings.h:298
    Settings& ccf_analysis(bool flag) { ccf_analysis_ = flag; return *this; }  // settings.h:305
  private:
    bool ccf_analysis_ = false;                                         // settings.h:320
    int seed_ = 0;                                                      // settings.h:324
    int num_trials_ = 1000;                                             // settings.h:325
};

}  // namespace mef::openpsa
