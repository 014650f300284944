"""Scheduler runtime: clock schedule, fence/compute executor, workers, timeline."""
