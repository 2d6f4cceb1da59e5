"""Stub (TEST INFRASTRUCTURE): cosyvoice/llm/llm.py:24 imports MulticlassAccuracy at module level; the flow encoder
classes are only reachable through cosyvoice/utils/class_utils.py, which imports the LLM module.  Never called."""
