"""``KMeans`` with the quantum knobs.

The reference keeps q-means in a separate class (``qMeans_``); its stock ``KMeans`` has no
``delta``.  This thin alias gives the name BASELINE.json uses: a ``KMeans`` whose
``delta`` / ``intermediate_error`` / ``true_distance_estimate`` arguments select the quantum
branch.  Unlike ``qMeans_`` the default ``algorithm='auto'`` resolves to the Lloyd path, since
that is the only one with a quantum branch."""
from ._dmeans import qMeans_


class KMeans(qMeans_):
    def __init__(self, n_clusters=8, *, init='k-means++', n_init=10, max_iter=300, tol=1e-4, verbose=0,
                 random_state=None, copy_x=True, algorithm='auto', delta=0, intermediate_error=False,
                 true_tomography=True, stop_when_reached_accuracy=True, true_distance_estimate=True):
        super().__init__(n_clusters=n_clusters, init=init, n_init=n_init, max_iter=max_iter, tol=tol, verbose=verbose,
                         random_state=random_state, copy_x=copy_x,
                         algorithm='full' if algorithm in ('auto', 'lloyd') else algorithm, delta=delta,
                         intermediate_error=intermediate_error, true_tomography=true_tomography,
                         stop_when_reached_accuracy=stop_when_reached_accuracy,
                         true_distance_estimate=true_distance_estimate)
