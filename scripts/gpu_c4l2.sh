#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
python scripts/c4_after_c3.py 2>&1 | tail -5
