// oracle/matchshim: empty stand-in for <boost/serialization/assume_abstract.hpp> (test infrastructure); names come from facade.h
