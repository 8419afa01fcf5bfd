;; UNVERIFIED: no JVM on the box this was written on; it has never been compiled.
;; `infer :pgibbs` (iterated conditional SMC) and `infer :pgas`.  Replaces
;; src/anglican/pgibbs.cljc:76-126: the retained particle sits in slot 0, the other N-1 slots are
;; resampled (pgibbs.cljc:92-99), log-weights are reset after the last conditional resample
;; (pgibbs.cljc:88), and the next retained path is the lineage of a uniformly chosen final
;; particle (`rand-nth`, pgibbs.cljc:122-123) traced back on the device (apop_trace_lineage =
;; `::past-samples`, pgibbs.cljc:29-35).  :pgas adds ancestor sampling at every conditional
;; observe (pgas.cljc:107-168).
(ns anglican.gpu.pgibbs
  (:require [anglican.inference :refer [infer]]
            [anglican.gpu.ffi :as ffi]
            [anglican.gpu.smc :as smc])
  (:import [java.lang.foreign Arena ValueLayout]))

(defn trace-lineage
  "x_0..x_T of final slot `slot` of the last keep-history sweep"
  [ctx slot T dim]
  (with-open [arena (Arena/ofConfined)]
    (let [out (.allocate arena ValueLayout/JAVA_DOUBLE (long (* (inc T) dim)))]
      (ffi/call ffi/apop-trace-lineage ctx (long slot) out)
      (vec (.toArray out ValueLayout/JAVA_DOUBLE)))))

(defn- csmc-seq [prog value number-of-particles device ancestor-sampling?]
  (let [{:keys [model observations]} (if (fn? prog) (prog value) prog)
        dim (:dim model 1)
        T (quot (count observations) dim)
        ctx (smc/create number-of-particles device (System/nanoTime))]
    (letfn [(sample-seq [retained i]
              (lazy-seq
                (.invokeWithArguments ffi/apop-set-seed [ctx (long i)])
                (let [out (smc/sweep ctx number-of-particles model observations
                                     :retained retained
                                     :csmc (if retained 1 0)
                                     ;; the last observe resamples unconditionally: pgibbs.cljc:89
                                     :n-cond-steps (if ancestor-sampling? T (dec T))
                                     :reset-weights 1
                                     :keep-history 2
                                     :ancestor-sampling (if (and retained ancestor-sampling?) 1 0))
                      slot (rand-int number-of-particles)            ; pgibbs.cljc:122
                      retained' (trace-lineage ctx slot T dim)]
                  (concat (smc/states->seq out dim)
                          (sample-seq retained' (inc i))))))]
      (sample-seq nil 0))))

(defmethod infer :pgibbs [_ prog value
                          & {:keys [number-of-particles device]
                             :or {number-of-particles 2 device 0}}]
  (assert (>= number-of-particles 2)
          ":number-of-particles must be at least 2")                 ; pgibbs.cljc:115-116
  (csmc-seq prog value number-of-particles device false))

(defmethod infer :pgas [_ prog value
                        & {:keys [number-of-particles device]
                           :or {number-of-particles 2 device 0}}]
  (assert (>= number-of-particles 2)
          ":number-of-particles must be at least 2")                 ; pgas.cljc:176-177
  (csmc-seq prog value number-of-particles device true))
