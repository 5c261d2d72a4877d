for g in 128 256; do for p in 0 1; do RAILS_PROFILE=$p python scripts/profile_sections.py C3 $g 2>&1 | tail -1; done; done
