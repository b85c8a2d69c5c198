"""Polling a TfJob to completion, the way the reference's Python client does.

Mirror of ``py/tf_job_client.py:55-96`` (``log_status``, ``wait_for_job``): same arguments,
same loop, same termination test (``results["status"]["phase"] == "Done"``) and the same
timeout rule.  Where the reference GETs the custom object from the API server
(``crd_api.get_namespaced_custom_object``), ``client`` here is the status file the local runner
rewrites after every reconcile (``python -m k8s_b200.run --workload syncps --status_file F``), or any
callable returning the TfJob object as a dict.
"""
from __future__ import annotations

import datetime
import json
import logging
import time

TF_JOB_GROUP = "tensorflow.org"
TF_JOB_VERSION = "v1alpha1"
TF_JOB_PLURAL = "tfjobs"
TF_JOB_KIND = "TfJob"


class TimeoutError(Exception):  # noqa: A001  (py/util.py:TimeoutError)
    """An error indicating an operation timed out."""


def log_status(tf_job):
    """A callback to use with wait_for_job (py/tf_job_client.py:55-61)."""
    logging.info("Job %s in namespace %s; phase=%s, state=%s,",
                 tf_job["metadata"]["name"],
                 tf_job["metadata"]["namespace"],
                 tf_job["status"]["phase"],
                 tf_job["status"]["state"])


def _get(client, namespace, name):
    if callable(client):
        return client(namespace, name)
    while True:                      # the runner replaces the file atomically; it may not exist yet
        try:
            with open(client) as f:
                return json.load(f)
        except (OSError, ValueError):
            time.sleep(0.02)


def wait_for_job(client, namespace, name,
                 timeout=datetime.timedelta(minutes=5),
                 polling_interval=datetime.timedelta(seconds=30),
                 status_callback=None):
    """Wait for the specified job to finish (py/tf_job_client.py:63-96)."""
    end_time = datetime.datetime.now() + timeout
    while True:
        results = _get(client, namespace, name)
        if results["metadata"]["name"] != name:
            raise KeyError("tfjobs %r not found in namespace %r" % (name, namespace))

        if status_callback:
            status_callback(results)

        if results["status"]["phase"] == "Done":
            return results

        if datetime.datetime.now() + polling_interval > end_time:
            raise TimeoutError(
                "Timeout waiting for job {0} in namespace {1} to finish.".format(
                    name, namespace))

        time.sleep(polling_interval.total_seconds())
