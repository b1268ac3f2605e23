// Kernel-selection / tuning overrides (apvae_tuning, include/apvae.h) as seen by the planner, the model builder and
// the runtime.  The options travel in apvae_config (model handles) or with apvae_plan_create (layer plans); the entry
// points that plan or build install them for the duration of the call (TuningScope), and the code that used to read
// APVAE_* environment variables reads tuning() instead.  Nothing in the library reads the environment.
#pragma once
#include "../../include/apvae.h"

namespace apvae {

const apvae_tuning& tuning();                     // the options of the call in progress on this thread (all-zero = defaults)

struct TuningScope {
    explicit TuningScope(const apvae_tuning* t);  // null: defaults
    ~TuningScope();
    TuningScope(const TuningScope&) = delete;
    TuningScope& operator=(const TuningScope&) = delete;
private:
    const apvae_tuning* prev_;
    apvae_tuning copy_;
};

// "switch" fields: 0 = the library's default, -1 = off, +1 = on
inline bool switch_on(int v, bool dflt) { return v == 0 ? dflt : v > 0; }

}  // namespace apvae
