;; UNVERIFIED: no JVM on the box this was written on; it has never been compiled.
;; `infer :pimh` - particle independent Metropolis-Hastings.  Replaces
;; src/anglican/pimh.cljc:39-71: every iteration is one GPU sweep, accepted against the
;; previous sweep's log Z by apop_pimh_accept (`accept?`, lmh.cljc:146-151 via pimh.cljc:64);
;; emitted states carry log-weight 0 (pimh.cljc:55-57).
(ns anglican.gpu.pimh
  (:require [anglican.inference :refer [infer]]
            [anglican.state :as st]
            [anglican.runtime :refer [sample* uniform-continuous]]
            [anglican.gpu.ffi :as ffi]
            [anglican.gpu.smc :as smc]))

(defmethod infer :pimh [_ prog value
                        & {:keys [number-of-particles device]
                           :or {number-of-particles 2 device 0}}]
  (assert (>= number-of-particles 1)
          ":number-of-particles must be at least 1")                 ; pimh.cljc:47-48
  (let [{:keys [model observations]} (if (fn? prog) (prog value) prog)
        dim (:dim model 1)
        ctx (smc/create number-of-particles device (System/nanoTime))
        run (fn [i]
              (.invokeWithArguments ffi/apop-set-seed [ctx (long i)])
              (smc/sweep ctx number-of-particles model observations))]
    (letfn [(add-samples [out i]
              (concat (map #(st/set-log-weight % 0.) (smc/states->seq out dim))
                      (sample-seq out (inc i))))
            (sample-seq [out i]
              (lazy-seq
                (let [next-out (run i)
                      u (sample* (uniform-continuous 0. 1.))]
                  (if (pos? (.invokeWithArguments ffi/apop-pimh-accept
                                                  [(double (:log-z next-out)) (double (:log-z out)) (double u)]))
                    (add-samples next-out i)
                    (add-samples out i)))))]
      (sample-seq (run 0) 1))))
