/* placeholder until the legacy env restatement lands */ typedef int usv_oracle_legacy_placeholder_t;
